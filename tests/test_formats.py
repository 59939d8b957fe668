"""SURVEY 8(f) "next" rows N1 / N2 on the CPU: the on-disk formats either side of the hot path -- `.gexf` graph
trees (src/data.py:67-83), `result/<dataset>/<metric>/*.npy` ground-truth matrices (src/results.py:182-189) -- and
the result dict `Eval.eval_test` hands to tuning.parse_results (eval.py:25-42, exp.py:337-369, 422-455)."""
import os

import networkx as nx
import numpy as np
import pytest

from util import corpus


def _write_tree(root, folder, train_gs, test_gs):
    for sub, gs in (('train', train_gs), ('test', test_gs)):
        d = os.path.join(root, folder, sub)
        os.makedirs(d)
        for g in gs:
            h = nx.Graph()
            for v, a in g.nodes(data=True):
                h.add_node(str(v), type=a['type'])
            for u, v in g.edges():
                h.add_edge(str(u), str(v), valence=1)
            nx.write_gexf(h, os.path.join(d, '{}.gexf'.format(g.graph['gid'])))


@pytest.fixture
def tree(tmp_path, monkeypatch):
    from graphembedding_b200 import data as D
    train_gs, test_gs = corpus(24, 3, 9, seed=11), corpus(6, 3, 9, seed=12)
    for i, g in enumerate(train_gs):
        g.graph['gid'] = 100 + i
    for i, g in enumerate(test_gs):
        g.graph['gid'] = 900 + i
    _write_tree(str(tmp_path / 'data'), 'AIDS700nef', train_gs, test_gs)
    monkeypatch.setattr(D, '_DATA_ROOT', str(tmp_path / 'data'))
    return tmp_path, train_gs, test_gs


def test_gexf_loader_reads_the_reference_layout(tree):
    """data/<folder>/{train,test}/<gid>.gexf, natural sort, gid from the file name, 'type' node attribute, valence
    stripped for the *nef sets (src/data.py:67-83, 119-121)."""
    import warnings
    from graphembedding_b200 import data as D
    tmp, train_gs, test_gs = tree
    with warnings.catch_warnings():
        warnings.simplefilter('error')            # real files: no synthetic-fallback warning
        tr, te = D.AIDS700nefData(True), D.AIDS700nefData(False)
    assert [g.graph['gid'] for g in tr.graphs] == [100 + i for i in range(24)]
    assert [g.graph['gid'] for g in te.graphs] == [900 + i for i in range(6)]
    for got, want in zip(tr.graphs, train_gs):
        assert got.number_of_nodes() == want.number_of_nodes() and got.number_of_edges() == want.number_of_edges()
        assert sorted(nx.get_node_attributes(got, 'type').values()) == sorted(nx.get_node_attributes(want, 'type').values())
        assert all('valence' not in ed for _, _, ed in got.edges(data=True))
    with pytest.warns(UserWarning, match='SYNTHETIC'):   # a missing tree falls back loudly
        D.AIDS10kNEFData(False)


def test_result_npy_loader_matches_reference_file_pattern(tree):
    from graphembedding_b200 import data as D
    from graphembedding_b200.dist_calculator import proxy_ged
    from graphembedding_b200.results import PairwiseGEDModelResult
    tmp, train_gs, test_gs = tree
    tr, te = D.load_data('aids700nef', True), D.load_data('aids700nef', False)
    dist = np.array([[proxy_ged(q, g) for g in tr.graphs] for q in te.graphs], np.float64)
    secs = np.full(dist.shape, 0.25)
    root = str(tmp / 'result')
    os.makedirs(root + '/aids700nef/ged')
    os.makedirs(root + '/aids700nef/time')
    np.save(root + '/aids700nef/ged/ged_ged_mat_aids700nef_astar_2018-01-01T00:00:00_host_10cpus.npy', dist)
    np.save(root + '/aids700nef/time/ged_time_mat_aids700nef_astar_2018-01-01T00:00:00_host_10cpus.npy', secs)
    r = PairwiseGEDModelResult.from_disk('aids700nef', 'astar', root=root)
    assert r.m_n() == (6, 24) and np.array_equal(r.dist_mat(False), dist) and np.array_equal(r.time_mat(), secs)
    qn = np.array([g.number_of_nodes() for g in te.graphs], np.float64)[:, None]
    dn = np.array([g.number_of_nodes() for g in tr.graphs], np.float64)[None, :]
    assert np.allclose(r.dist_mat(True), 2 * dist / (qn + dn), rtol=0, atol=0)       # distance.py:59-60
    assert np.array_equal(r.sort_id_mat(False), np.argsort(dist, kind='mergesort'))
    with pytest.raises(RuntimeError, match='No results found'):
        PairwiseGEDModelResult.from_disk('aids700nef', 'beam80', root=root)


def test_eval_test_result_dict_has_the_reference_keys_and_shapes(tree):
    """eval.py:25-42 -> exp.py plot_apk (:337-369: ks = 1 .. n-1, per model {'ks', 'aps'}) and plot_mrr_mse_time
    (:422-455: per model scalar) for norm in (True, False): exactly what tuning.parse_results reads (tuning.py:175-184)."""
    from graphembedding_b200 import data as D
    from graphembedding_b200 import metrics
    from graphembedding_b200.dist_calculator import proxy_ged
    from graphembedding_b200.eval import Eval
    from graphembedding_b200.results import PairwiseGEDModelResult
    metrics.use_engine(None) if hasattr(metrics, 'use_engine') else None
    tmp, train_gs, test_gs = tree
    tr, te = D.load_data('aids700nef', True), D.load_data('aids700nef', False)
    dist = np.array([[proxy_ged(q, g) for g in tr.graphs] for q in te.graphs], np.float64)
    truth = PairwiseGEDModelResult('aids700nef', 'astar', dist, [g.number_of_nodes() for g in te.graphs],
                                   [g.number_of_nodes() for g in tr.graphs], time_mat=np.zeros(dist.shape))
    ev = Eval('aids700nef', 'astar', 'gaussian', 0.6, true_result=truth)
    rng = np.random.default_rng(0)
    sim = np.clip(truth.sim_mat('gaussian', 0.6, True) + 0.05 * rng.standard_normal(dist.shape), 0, 1)
    out = ev.eval_test('siamese_gcntn_mse', sim, np.full(dist.shape, 1e-3))
    m, n = dist.shape
    want = {'apk_norm', 'apk_nonorm', 'mrr_norm', 'mrr_nonorm', 'mse_norm', 'mse_nonorm', 'time_norm', 'time_nonorm'}
    assert want <= set(out)
    assert set(out) - want <= {k for k in out if k.startswith('kendall_tau')}          # additive keys only
    for k in ('apk_norm', 'apk_nonorm'):
        assert set(out[k]) == {'astar', 'siamese_gcntn_mse'}
        for mdl, rec in out[k].items():
            assert list(rec['ks']) == list(range(1, n)) and len(rec['aps']) == n - 1
            assert all(0.0 <= a <= 1.0 for a in rec['aps'])
        assert all(a == 1.0 for a in out[k]['astar']['aps'])                              # the truth ranks itself perfectly
    for k in want - {'apk_norm', 'apk_nonorm'}:
        assert set(out[k]) == {'astar', 'siamese_gcntn_mse'} and all(np.isscalar(v) or np.ndim(v) == 0 for v in out[k].values())
    assert out['mrr_norm']['astar'] == 1.0 and out['mse_norm']['astar'] == 0.0
