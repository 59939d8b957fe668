"""Pin the CPU oracle (oracle/siamese_oracle.py) against golden vectors produced by RUNNING
THE REFERENCE in the build container (tests/golden/make_golden.py): the reference's own host
modules, and its unmodified model code executed on oracle/tf1_shim.py."""
import json
import os

import numpy as np
import pytest

from util import O

torch = pytest.importorskip('torch')
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')


def _load(name):
    with open(os.path.join(G, name)) as f:
        return json.load(f)


def _oracle_graph(rec, feat_idx):
    return O.Graph.from_edges([feat_idx[t] for t in rec['types']], [tuple(e) for e in rec['edges']])


def test_preprocessing_matches_reference_modelgraph():
    d = _load('preprocess.json')
    for g, ref in zip(d['graphs'], d['ref']):
        n = len(g['types'])
        rows, cols, vals = O.normalized_adjacency_coo(n, [tuple(e) for e in g['edges']])
        lc = np.array(ref['lap_coords']).reshape(-1, 2)
        lv = np.array([float.fromhex(h) for h in ref['lap_vals_hex']])
        assert np.array_equal(np.stack([rows, cols], 1), lc)          # same COO order
        assert np.array_equal(vals, lv)                               # bit-exact float64
        assert ref['lap_shape'] == [n, n]
        # features: one-hot, row-normalised (graphs.py:55-62)
        fc = np.array(ref['feat_coords']).reshape(-1, 2)
        assert np.array_equal(fc[:, 0], np.arange(n))
        assert np.array_equal(fc[:, 1], [d['feat_idx_dic'][t] for t in g['types']])
        assert ref['feat_vals'] == [1.0] * n and ref['nnz_shape'] == [n]


def test_random_sampler_sequence():
    d = _load('samplers.json')
    for key, seq in d.items():
        if not key.startswith('random_'):
            continue
        L = int(key.split('_')[1])
        s = O.RandomSamplerOracle(list(range(L)))
        assert [list(s.get_pair()) for _ in range(len(seq))] == seq


def test_similarity_kernel_and_normalized_dist():
    d = _load('similarity.json')
    ds = np.array(d['d'])
    for y in (0.6, 0.3, 0.2, 0.01):
        ref = np.array([float.fromhex(h) for h in d['gauss_%g' % y]])
        assert np.array_equal(O.gaussian_sim(ds, y), ref)
    for dist, n1, n2, h in d['normalized_dist']:
        assert O.normalized_dist(dist, n1, n2) == float.fromhex(h)


def test_ranking_and_metrics_match_reference():
    d = _load('metrics.json')
    dist, norm, pred = np.array(d['dist']), np.array(d['norm']), np.array(d['pred'])
    ks, m, n = d['ks'], d['m'], d['n']
    ps = O.sort_id_mat_sim(pred)
    assert np.array_equal(ps, np.array(d['pred_sort_id']))
    for key, tm in (('norm', norm), ('nonorm', dist)):
        ts = O.sort_id_mat_dist(tm)
        assert np.array_equal(ts, np.array(d['true_sort_id_' + key]))
        assert np.allclose(O.precision_at_ks(ts, tm, ps, pred, ks), d['prec_' + key], rtol=0, atol=1e-15)
        assert np.allclose(O.precision_at_ks(ts, tm, ps, pred, list(range(1, n))), d['prec_all_' + key],
                           rtol=0, atol=1e-15)
        assert abs(O.mean_reciprocal_rank(ts, tm, ps, pred) - d['mrr_' + key]) < 1e-15
        assert abs(O.frobenius_mse(O.gaussian_sim(tm, 0.6), pred) - d['mse_' + key]) < 1e-15
        for i in range(m):
            for ki, k in enumerate(ks):
                assert O.top_k_ids(ts[i], tm[i], k, True).tolist() == d['true_topk_incl_' + key][i][ki]
            assert [O.ranking(ts[i], tm[i], j) for j in range(n)] == d['true_rank_' + key][i]
    for i in range(m):
        for ki, k in enumerate(ks):
            assert O.top_k_ids(ps[i], pred[i], k, False).tolist() == d['pred_topk'][i][ki]
        assert [O.ranking(ps[i], pred[i], j) for j in range(n)] == d['pred_rank'][i]


def test_kendall_tau_matches_scipy():
    from scipy.stats import kendalltau
    rng = np.random.default_rng(0)
    for _ in range(20):
        x = rng.integers(0, 6, size=40).astype(float)
        y = np.round(rng.uniform(0, 1, size=40), 1)
        ref = kendalltau(x, y).statistic
        assert abs(O.kendall_tau_b(x, y) - ref) < 1e-12


MODELS = ['default', 'avg_ntn', 'dot']


def _model_setup(name):
    d = _load('model_%s.json' % name)
    fl = d['flags']
    specs = [fl['layer_%d' % i] for i in range(fl['num_layers'])]
    layers = O.parse_layer_specs(specs, d['input_dim'])
    # variable creation order of the reference == our flat blob order
    tab = O.param_table(layers)
    assert [(v[1], tuple(v[2])) for v in d['var_names']] == [(nm, tuple(sh)) for _, nm, sh in tab]
    graphs = {g['gid']: _oracle_graph(g, d['feat_idx_dic']) for g in d['train_graphs'] + d['test_graphs']}
    nodes = {g['gid']: len(g['types']) for g in d['train_graphs'] + d['test_graphs']}
    return d, fl, layers, graphs, nodes


@pytest.mark.parametrize('name', MODELS)
def test_forward_matches_reference_model_code(name):
    """test() loop of the reference (train.py:57-61, 98-101) vs the restatement."""
    d, fl, layers, graphs, _ = _model_setup(name)
    P = O.unflatten(layers, np.array(d['params0']), torch.float64)
    raw_ref, sim_ref = np.array(d['test_raw']), np.array(d['test_sim'])
    for i, qg in enumerate(d['q_gids']):
        x1 = O.graph_forward(layers, P, graphs[qg], d['input_dim'])
        for j, dg in enumerate(d['db_gids']):
            x2 = O.graph_forward(layers, P, graphs[dg], d['input_dim'])
            r = float(O.pair_forward(layers, P, x1, x2, 'reference'))
            assert abs(r - raw_ref[i, j]) <= 1e-12 * max(1.0, abs(raw_ref[i, j]))
            s = float(O._act(fl['final_act'], fl['yeta'])(torch.tensor(r, dtype=torch.float64)))
            assert abs(s - sim_ref[i, j]) <= 1e-12


@pytest.mark.parametrize('name', MODELS)
def test_train_steps_match_reference_model_code(name):
    """loss (incl. F5/F6 broadcasts and all-variable weight decay), autodiff gradients and
    TF-Adam updates of the reference (models.py:28-36, 67-88; model_mse.py:145-151) vs the
    restatement, fed with the pairs/labels the reference's get_feed_dict drew (F7)."""
    d, fl, layers, graphs, nodes = _model_setup(name)
    B = fl['batch_size']
    adam = O.TFAdam(len(d['params0']), lr=fl['learning_rate'])
    p = np.array(d['params0'])
    for st in d['steps']:
        assert np.allclose(p, st['params_before'], rtol=0, atol=1e-12)
        calls = st['sampler_calls']
        assert len(calls) == B + B * B                       # F7: B inputs + B*B label draws
        g1 = [graphs[c[1]] for c in calls[:B]]
        g2 = [graphs[c[2]] for c in calls[:B]]
        lab = calls[-B:]                                     # the labels that survive: last B draws
        dist = st['norm_dists'] if fl['dist_norm'] else st['dists']
        from graphembedding_b200.dist_calculator import proxy_ged  # noqa: F401  (same proxy the golden used)
        tgt = O.gaussian_sim(np.array(dist), fl['yeta'])
        # normalised distance restated (distance.py:59-60) agrees with what the reference fed
        for (tv, a, b), dd, nd in zip(lab, st['dists'], st['norm_dists']):
            assert abs(O.normalized_dist(dd, nodes[a], nodes[b]) - nd) < 1e-15
        loss, mse, pred, g, raw = O.loss_and_grads(layers, p, g1, g2, tgt, d['input_dim'],
                                                   final_act=fl['final_act'], yeta=fl['yeta'],
                                                   weight_decay=fl['weight_decay'])
        assert abs(loss - st['loss']) <= 1e-12 * max(1.0, abs(st['loss']))
        assert np.allclose(g, st['grads'], rtol=1e-9, atol=1e-13)
        p = adam.step(p, g)
        assert np.allclose(p, st['params_after'], rtol=0, atol=1e-11)


def _attention_case(rec):
    """A model whose node rows ARE the fixture's x: n isolated nodes with distinct labels (A_hat = I), one identity
    GraphConvolution without bias whose weight matrix is x, then Attention(d)."""
    d, n = rec['d'], rec['n']
    specs = ['GraphConvolution:output_dim=%d,act=identity,dropout=False,bias=False,sparse_inputs=True' % d,
             'Attention:input_dim=%d' % d, 'Dot']
    layers = O.parse_layer_specs(specs, n)
    g = O.Graph.from_edges(list(range(n)), [])
    return specs, layers, g, np.array(rec['x']), np.array(rec['weights']), np.array(rec['out'])


def test_attention_pooling_is_pinned_to_the_reference_class():
    """layers.py:143-160 run by make_golden.py (reference class on the TF-1 stand-in): the oracle's Attention branch,
    per graph and batched, must reproduce it (the factory cannot reach the class, SURVEY F2, but its math is the pin)."""
    import torch
    from oracle import batched as OB
    recs = json.load(open(os.path.join(G, 'attention.json')))
    assert len(recs) >= 4
    for rec in recs:
        specs, layers, g, x, W, out = _attention_case(rec)
        P = {(0, 'weights_0'): torch.as_tensor(x), (1, 'weights'): torch.as_tensor(W)}
        got = O.graph_forward(layers, P, g, rec['n'], torch.float64).numpy()
        assert np.max(np.abs(got - out)) <= 1e-12
        lab, A, mask, n = OB.pad_graphs([g], rec['n'], torch.float64)
        got_b = OB.embed(layers, P, lab, A, mask, n, torch.float64).numpy()[0]
        assert np.max(np.abs(got_b - out)) <= 1e-12
