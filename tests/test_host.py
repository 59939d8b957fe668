"""CPU tests of the host side: packer, samplers, similarity, results/metrics, layer grammar,
C-ABI surface -- each against golden vectors written by the reference (tests/golden)."""
import json
import os
import re
import types

import numpy as np
import pytest
import networkx as nx

from util import O, ROOT

G = os.path.join(ROOT, 'tests', 'golden')


def _load(name):
    with open(os.path.join(G, name)) as f:
        return json.load(f)


def _nx(rec):
    g = nx.Graph()
    for i, t in enumerate(rec['types']):
        g.add_node(str(i), type=t)
    for u, v in rec['edges']:
        g.add_edge(str(u), str(v))
    g.graph['gid'] = rec['gid']
    return g


def test_c_abi_exports_every_declared_symbol():
    from graphembedding_b200 import native
    hdr = open(os.path.join(ROOT, 'include', 'sgnn.h')).read()
    declared = set(re.findall(r'\b(sgnn_[a-z_0-9]+)\s*\(', hdr))
    assert declared == set(native.SIGNATURES), declared ^ set(native.SIGNATURES)
    lib = native.lib()
    for name in declared:
        assert getattr(lib, name) is not None
    assert lib.sgnn_abi_version() == native.ABI_VERSION


def test_create_without_gpu_fails_loudly():
    torch = pytest.importorskip('torch')
    if torch.cuda.is_available():
        pytest.skip('GPU present')
    from graphembedding_b200.engine import SgnnEngine
    from graphembedding_b200.layers_factory import parse_layers
    from graphembedding_b200.native import SgnnError
    with pytest.raises(SgnnError):
        SgnnEngine(parse_layers(O.DEFAULT_LAYERS, 29), 29)


def test_packer_is_bit_exact_against_reference_modelgraph():
    from graphembedding_b200.graphs import ModelGraph, ModelGraphList, NodeFeatureOneHotEncoder
    d = _load('preprocess.json')
    gs = [_nx(r) for r in d['graphs']]
    enc = NodeFeatureOneHotEncoder(gs, 'type', feat_idx_dic=d['feat_idx_dic'])
    assert enc.input_dim() == d['input_dim']
    lst = ModelGraphList('random', -1, False, gs, enc)
    for mg, ref in zip(lst.gs, d['ref']):
        fc, fv, fs = mg.get_node_inputs()
        assert fc.tolist() == ref['feat_coords'] and fv.tolist() == ref['feat_vals']
        assert list(fs) == ref['feat_shape'] and list(mg.get_node_inputs_num_nonzero()) == ref['nnz_shape']
        lc, lv, ls = mg.get_laplacians()[0]
        assert lc.tolist() == ref['lap_coords'] and list(ls) == ref['lap_shape']
        ref32 = np.array([float.fromhex(h) for h in ref['lap_vals_hex']]).astype(np.float32)
        assert np.array_equal(lv.astype(np.float32).view(np.uint32), ref32.view(np.uint32))
    # one graph packed alone == the same graph inside a batch
    solo = ModelGraph(gs[5], enc)
    assert np.array_equal(solo.csr()[2], lst.gs[5].csr()[2])


def test_packer_float32_rounding_for_all_degree_pairs():
    """float32((pow(d_i,-.5))*(pow(d_j,-.5))) from the C++ packer == numpy's, for star graphs
    covering degrees 1..64."""
    from graphembedding_b200.native import PackedBatch
    labels, edges = [], []
    for deg in range(1, 65):
        labels.append(np.zeros(deg + 1, np.int32))
        edges.append(np.array([(0, i) for i in range(1, deg + 1)], np.int32))
    b = PackedBatch.from_lists(labels, edges)
    for g, deg in enumerate(range(1, 65)):
        r, c, v = O.normalized_adjacency_coo(deg + 1, [(0, i) for i in range(1, deg + 1)])
        rp, cc, vv = O.coo_to_canonical_csr(deg + 1, r, c, v)
        lo = b.rowptr[b.node_off[g]]; hi = b.rowptr[b.node_off[g + 1]]
        assert np.array_equal(b.val[lo:hi].view(np.uint32), vv.view(np.uint32))
        assert np.array_equal(b.col[lo:hi], cc)


def test_packer_edge_cases():
    from graphembedding_b200.native import PackedBatch
    b = PackedBatch.from_lists([], [])
    assert b.n_graphs == 0 and b.nnz == 0
    # single node, duplicate edges, self loop
    b = PackedBatch.from_lists([[0], [1, 2, 0]], [np.zeros((0, 2), np.int32), [(0, 1), (1, 0), (2, 2)]])
    assert b.n_graphs == 2 and b.max_nodes == 3
    assert b.val[0] == 1.0                                     # isolated node: 1/sqrt(1)^2
    sub = b.select([1])
    assert sub.n_graphs == 1 and sub.total_nodes == 3
    cat = PackedBatch.concat([sub, b])
    assert cat.n_graphs == 3 and cat.nnz == sub.nnz + b.nnz


def test_samplers_reproduce_reference_sequences():
    from graphembedding_b200.samplers import DistributionSampler, RandomSampler
    from graphembedding_b200 import data as D
    d = _load('samplers.json')
    gs = D.synthetic_corpus(40, 1, 10, seed=501) + D.synthetic_corpus(8, 11, 30, seed=502, gid_base=100)

    class It(object):
        def __init__(self, i, g):
            self.i, self.g = i, g

        def get_nxgraph(self):
            return self.g
    for key, seq in d.items():
        parts = key.split('_')
        if parts[0] == 'random':
            L = int(parts[1])
            s = RandomSampler([It(i, gs[i % len(gs)]) for i in range(L)], -1, False)
        else:
            L, sn = int(parts[1]), int(parts[2])
            s = DistributionSampler([It(i, gs[i]) for i in range(L)], sn, False)
        assert [[a.i, b.i] for a, b in (s.get_pair() for _ in range(len(seq)))] == seq, key


def test_similarity_and_distance_interfaces():
    from graphembedding_b200.similarity import create_sim_kernel
    from graphembedding_b200.distance import normalized_dist
    d = _load('similarity.json')
    ds = np.array(d['d'])
    for y in (0.6, 0.3, 0.2, 0.01):
        ref = np.array([float.fromhex(h) for h in d['gauss_%g' % y]])
        assert np.array_equal(create_sim_kernel('gaussian', y).dist_to_sim_np(ds), ref)
    assert np.array_equal(create_sim_kernel('identity').dist_to_sim_np(ds), np.array(d['identity']))
    with pytest.raises(RuntimeError):
        create_sim_kernel('linear')
    for dist, n1, n2, h in d['normalized_dist']:
        g1, g2 = nx.path_graph(n1), nx.path_graph(n2)
        assert normalized_dist(dist, g1, g2) == float.fromhex(h)


def test_results_and_metrics_match_reference():
    from graphembedding_b200 import metrics as MT
    from graphembedding_b200.results import PairwiseGEDModelResult, SiameseModelResult
    d = _load('metrics.json')
    m, n, ks = d['m'], d['n'], d['ks']
    tr = PairwiseGEDModelResult('syn', 'astar', np.array(d['dist']), d['q_nodes'], d['db_nodes'])
    pr = SiameseModelResult('syn', 'siamese_gcntn_mse', sim_mat=np.array(d['pred']), time_mat=np.ones((m, n)))
    assert np.array_equal(tr.dist_mat(True), np.array(d['norm']))
    assert np.array_equal(pr.sort_id_mat(), np.array(d['pred_sort_id']))
    for nm, key in ((True, 'norm'), (False, 'nonorm')):
        assert np.array_equal(tr.sort_id_mat(nm), np.array(d['true_sort_id_' + key]))
        assert np.allclose(MT.precision_at_ks(tr, pr, nm, ks), d['prec_' + key], rtol=0, atol=1e-15)
        assert np.allclose(MT.precision_at_ks(tr, pr, nm, list(range(1, n))), d['prec_all_' + key], rtol=0, atol=1e-15)
        assert abs(MT.mean_reciprocal_rank(tr, pr, nm) - d['mrr_' + key]) < 1e-15
        assert abs(MT.mean_squared_error(tr, pr, 'gaussian', 0.6, nm) - d['mse_' + key]) < 1e-15
        for i in range(m):
            for ki, k in enumerate(ks):
                assert tr.top_k_ids(i, k, nm, True).tolist() == d['true_topk_incl_' + key][i][ki]
            assert [tr.ranking(i, j, nm) for j in range(n)] == d['true_rank_' + key][i]
    for i in range(m):
        for ki, k in enumerate(ks):
            assert pr.top_k_ids(i, k, False, False).tolist() == d['pred_topk'][i][ki]
        assert [pr.ranking(i, j, False) for j in range(n)] == d['pred_rank'][i]
    assert MT.average_time(pr) == d['time']
    with pytest.raises(RuntimeError):
        pr.top_k_ids(0, n, False, False)                       # results.py:68-69


def test_kendall_tau_counts_exact():
    from scipy.stats import kendalltau
    from graphembedding_b200 import metrics as MT
    rng = np.random.default_rng(3)
    for n in (2, 3, 17, 200):
        x = rng.integers(0, 5, size=n).astype(float)
        y = np.round(rng.uniform(0, 1, size=n), 1)
        got = MT.kendall_tau_counts(x, y)
        assert got == O.kendall_tau_counts(x, y)
        con, dis, tx, ty, txy = got
        den = np.sqrt(float(con + dis + tx)) * np.sqrt(float(con + dis + ty))
        if den > 0:
            assert abs((con - dis) / den - kendalltau(x, y).statistic) < 1e-12


def test_layer_grammar_and_errors():
    from graphembedding_b200.layers_factory import parse_layers, parse_as_bool
    assert parse_layers(O.DEFAULT_LAYERS, 29) == O.parse_layer_specs(O.DEFAULT_LAYERS, 29)
    for bad, msg in (('Foo:a=1', 'Unknown layer Foo'), ('Average:x=1', 'Average layer must have 0 specs'),
                     ('Padding:max_in_dims=10', 'Padding layer must have 2 specs'),
                     ('Dot:x=1', 'Dot layer must have 0 specs')):
        with pytest.raises(RuntimeError, match=msg):
            parse_layers([bad], 29)
    with pytest.raises(RuntimeError, match='Unknown bool string'):
        parse_as_bool('true')
    with pytest.raises(RuntimeError, match='must be specified'):
        parse_layers([O.DEFAULT_LAYERS[0], O.DEFAULT_LAYERS[0]], 29)
    with pytest.raises(RuntimeError, match='Unknown activation'):
        parse_layers(['Dense:input_dim=16,output_dim=1,dropout=True,act=gelu,bias=True'], 29)


def test_flag_defaults_match_reference_config():
    from graphembedding_b200.config import ensure_parsed
    F = ensure_parsed()
    ref = dict(dataset='aids80nef', valid_percentage=0.25, node_feat_name='type', node_feat_encoder='onehot',
               edge_feat_name='valence', edge_feat_processor=None, dist_metric='ged', dist_algo='astar',
               sampler='random', sample_num=-1, sampler_duplicate_removal=False, model='siamese_gcntn_mse',
               num_layers=5, batch_size=5, dist_norm=True, sim_kernel='gaussian', yeta=0.6,
               final_act='sim_kernel', loss_func='mse', delta=0.1, gamma=0.1, num_neg=5, graph_loss=None,
               dropout=0.1, weight_decay=5e-4, learning_rate=0.01, iters=20, early_stopping=None, log=True,
               plot_results=False)
    d = F.flag_values_dict()
    for k, v in ref.items():
        assert d[k] == v, k
    assert [d['layer_%d' % i] for i in range(5)] == O.DEFAULT_LAYERS


@pytest.mark.parametrize('name', ['default', 'avg_ntn'])
def test_feed_construction_draws_like_the_reference(name):
    """get_feed_dict advances the sampler B + B*B times and labels with the LAST B draws
    (model_mse.py:58-59, 82-92), reproducing the gids the reference fed."""
    from graphembedding_b200.data_siamese import SiameseModelData
    from graphembedding_b200.dist_calculator import DistCalculator
    from graphembedding_b200.model_mse import SiameseGCNTNMSE
    d = _load('model_%s.json' % name)
    fl = d['flags']
    tr, te = [_nx(r) for r in d['train_graphs']], [_nx(r) for r in d['test_graphs']]
    loader = lambda ds, train: types.SimpleNamespace(graphs=list(tr) if train else list(te))  # noqa: E731
    data = SiameseModelData(loader=loader, feat_idx_dic=d['feat_idx_dic'], valid_percentage=fl['valid_percentage'])
    assert data.input_dim() == d['input_dim']
    dc = DistCalculator('aids80nef', 'ged', 'astar')
    fake = types.SimpleNamespace(batch_size=fl['batch_size'], label_mode='reference', _offsets=SiameseGCNTNMSE._offsets,
                                 _gid=lambda off, mg: off[id(mg.batch)] + mg.index)
    all_graphs = tr + te
    order = [g.graph['gid'] for g in tr[:data.train_data.num_graphs()]] + \
            [g.graph['gid'] for g in tr[data.train_data.num_graphs():]] + [g.graph['gid'] for g in te]
    for st in d['steps']:
        for tvt, calls, dd, nd in (('train', st['sampler_calls'], st['dists'], st['norm_dists']),
                                   ('val', st['val_sampler_calls'], st['val_dists'], st['val_norm_dists'])):
            feed = SiameseGCNTNMSE.get_feed_dict(fake, data, dc, tvt, None, None)
            B = fl['batch_size']
            assert [order[i] for i in feed.g1] == [c[1] for c in calls[:B]]
            assert [order[i] for i in feed.g2] == [c[2] for c in calls[:B]]
            assert feed.dists.ravel().tolist() == dd
            assert feed.norm_dists.ravel().tolist() == nd


@pytest.mark.parametrize('n_hi,num', [(10, 300), (30, 64), (200, 6)])
def test_compact_wire_format_round_trip_is_bit_exact(n_hi, num):
    """sgnn_compact_batch -> sgnn_expand_compact_host (the host restatement of the device expansion kernel)
    must reproduce the five arrays of the batch bit for bit; the blob is ~3.5x smaller for AIDS-sized graphs."""
    from graphembedding_b200.native import PackedBatch
    from util import corpus, encoder_for
    from graphembedding_b200.graphs import pack_nx_graphs
    gs = corpus(num, 2, n_hi, seed=17)
    b = pack_nx_graphs(gs, encoder_for(gs))
    blob = b.compact(pinned=False)
    assert blob is not None and blob.numel() % 16 == 0
    back = PackedBatch.from_compact(blob)
    for f in PackedBatch.FIELDS:
        x, y = np.asarray(getattr(b, f)), np.asarray(getattr(back, f))
        assert x.dtype == y.dtype and x.shape == y.shape, f
        assert x.tobytes() == y.tobytes(), f
    assert (back.n_graphs, back.total_nodes, back.nnz, back.max_nodes) == (b.n_graphs, b.total_nodes, b.nnz, b.max_nodes)
    if n_hi == 10:
        assert blob.numel() < b.nbytes() / 3


def test_compact_wire_format_rejects_what_does_not_fit_a_byte():
    from graphembedding_b200.native import PackedBatch
    # a 300-node path graph: node indices exceed the byte-wide column field
    n = 300
    b = PackedBatch.from_lists([np.zeros(n, np.int32)], [np.array([[i, i + 1] for i in range(n - 1)], np.int32)])
    assert b.compact(pinned=False) is None
    # label 256 does not fit either
    b = PackedBatch.from_lists([np.array([0, 256], np.int32)], [np.array([[0, 1]], np.int32)])
    assert b.compact(pinned=False) is None
    # empty batch and single isolated node are fine
    e = PackedBatch.from_lists([], [])
    blob = e.compact(pinned=False)
    assert blob is not None and PackedBatch.from_compact(blob).n_graphs == 0
    one = PackedBatch.from_lists([np.array([3], np.int32)], [np.zeros((0, 2), np.int32)])
    back = PackedBatch.from_compact(one.compact(pinned=False))
    assert back.val.tolist() == [1.0] and back.labels.tolist() == [3] and back.rowptr.tolist() == [0, 1]


def test_tuning_tables_and_csv_layout_match_the_reference():
    """the sweep driver's tables are the reference's (tests/golden/tuning.json, extracted from
    model/Siamese/tuning.py with tests/golden/make_tuning_golden.py) and one sweep point writes the reference's
    CSV layout (tuning.py:113-158)."""
    from graphembedding_b200 import tuning
    g = _load('tuning.json')
    for name in ('dataset', 'header', 'model', 'dist_norm_range', 'yeta_range_norm', 'yeta_range_nonorm',
                 'final_act_range', 'lr_range', 'iter_range', 'val_ratio_range', 'dropout_range', 'layers'):
        assert getattr(tuning, name) == g[name], name
    pts = list(tuning.grid_points())
    assert len(pts) == (6 + 4) * 5 * 6 * 4 * 2 * 2 and pts[0] == (True, 0.1, 'identity', 0.06, 500, 0.25, 0.1)
    assert pts[-1] == (False, 0.0005, 'sim_kernel', 0.001, 2000, 0.4, 0.5)
    aps = np.linspace(0.1, 1.0, 12)
    res = {k: {tuning.model: v} for k, v in (('mrr_norm', 0.5), ('mrr_nonorm', 0.4), ('mse_norm', 0.01),
                                              ('mse_nonorm', 0.02))}
    res['apk_norm'] = {tuning.model: {'aps': aps}}
    res['apk_nonorm'] = {tuning.model: {'aps': aps[::-1]}}
    row = tuning.parse_results(res)
    assert len(row) == 24 and row[:10] == list(aps[:10]) and row[-4:] == [0.5, 0.4, 0.01, 0.02]
    import tempfile
    calls = []

    def fake_main():
        from graphembedding_b200.config import FLAGS
        calls.append((FLAGS.dist_norm, FLAGS.yeta, FLAGS.final_act, FLAGS.learning_rate, FLAGS.iters,
                      FLAGS.valid_percentage, FLAGS.dropout, FLAGS.num_layers, FLAGS.layer_2))
        return [0.3, 0.1 + 0.01 * len(calls), 0.2], [0.0] * 3, [0.5, 0.4 - 0.01 * len(calls)], [0.0] * 2, res

    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, 'sweep.csv')
        best_train, best_val = tuning.tune(file_name=path, limit=3, run=fake_main)
        lines = open(path).read().strip().split('\n')
    assert calls[0] == (True, 0.1, 'identity', 0.06, 500, 0.25, 0.1, 4, 'Average') and len(calls) == 3
    assert lines[0].split(',') == tuning.header
    assert len(lines) == 1 + 3 + 5 and lines[4] == 'Final results:'
    assert len(lines[1].split(',')) == len(tuning.header)
    assert best_train[6] == pytest.approx(0.11) and best_val[6] == pytest.approx(0.37)
    assert len(lines[5].split(',')) == len(lines[6].split(',')) == len(tuning.header) - 2


def test_header_constants_match_the_python_binding():
    """Enumerations of include/sgnn.h that the Python side mirrors by value."""
    from graphembedding_b200 import native
    from graphembedding_b200.engine import SgnnEngine
    hdr = open(os.path.join(ROOT, 'include', 'sgnn.h')).read()

    def const(name):
        m = re.search(r'\b' + name + r'\s*=\s*(\d+)', hdr) or re.search(r'#define\s+' + name + r'\s+(\d+)', hdr)
        assert m, name
        return int(m.group(1))
    for kind, v in native.LAYER_KINDS.items():
        cname = {'GraphConvolution': 'SGNN_LAYER_GCN'}.get(kind, 'SGNN_LAYER_' + kind.upper())
        assert const(cname) == v, kind
    for act, v in native.ACTS.items():
        assert const('SGNN_ACT_' + act.upper()) == v, act
    assert const('SGNN_PREC_FP32') == native.PRECISIONS['fp32'] and const('SGNN_PREC_BF16X3') == native.PRECISIONS['bf16x3']
    assert const('SGNN_STAGE_PREP_QUERIES') == SgnnEngine.STAGE_PREP_QUERIES
    assert const('SGNN_STAGE_PREP_DATABASE') == SgnnEngine.STAGE_PREP_DATABASE
    assert const('SGNN_STAGE_SCORE') == SgnnEngine.STAGE_SCORE
    assert const('SGNN_MAX_LAYERS') == native.SGNN_MAX_LAYERS and const('SGNN_ABI_VERSION') == native.ABI_VERSION
