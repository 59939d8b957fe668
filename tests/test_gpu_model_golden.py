"""End to end on the GPU through the reference-shaped host API (SiameseModelData, create_model,
run_tf, train_val-style stepping, all-pairs test) against the golden run of the reference's
own model code (tests/golden/model_*.json)."""
import json
import os
import types

import networkx as nx
import numpy as np
import pytest

from util import ROOT, close_fp32

torch = pytest.importorskip('torch')
pytestmark = pytest.mark.gpu
G = os.path.join(ROOT, 'tests', 'golden')


def _nx(rec):
    g = nx.Graph()
    for i, t in enumerate(rec['types']):
        g.add_node(str(i), type=t)
    for u, v in rec['edges']:
        g.add_edge(str(u), str(v))
    g.graph['gid'] = rec['gid']
    return g


def _setup(name):
    from graphembedding_b200.data_siamese import SiameseModelData
    from graphembedding_b200.dist_calculator import DistCalculator
    from graphembedding_b200.train import Saver, Session, create_model
    with open(os.path.join(G, 'model_%s.json' % name)) as f:
        d = json.load(f)
    fl = d['flags']
    tr, te = [_nx(r) for r in d['train_graphs']], [_nx(r) for r in d['test_graphs']]
    loader = lambda ds, train: types.SimpleNamespace(graphs=list(tr) if train else list(te))  # noqa: E731
    data = SiameseModelData(loader=loader, feat_idx_dic=d['feat_idx_dic'], valid_percentage=fl['valid_percentage'])
    specs = [fl['layer_%d' % i] for i in range(fl['num_layers'])]
    model = create_model('siamese_gcntn_mse', data.input_dim(), layer_specs=specs, batch_size=fl['batch_size'],
                         dropout=0.0, yeta=fl['yeta'], final_act=fl['final_act'], dist_norm=fl['dist_norm'],
                         weight_decay=fl['weight_decay'], learning_rate=fl['learning_rate'])
    model.engine.set_params(np.array(d['params0'], np.float32))
    sess = Session(model)
    return d, fl, data, DistCalculator('aids80nef', 'ged', 'astar'), model, sess, Saver(sess)


@pytest.mark.parametrize('name', ['default', 'avg_ntn', 'dot'])
def test_all_pairs_test_matrix_matches_reference_run(name):
    from graphembedding_b200.train import all_pairs_similarity, run_tf
    d, fl, data, dc, model, sess, saver = _setup(name)
    ref = np.array(d['test_sim'])
    sim, ids, _ = all_pairs_similarity(data, model, k=3)
    assert sim.shape == ref.shape
    assert close_fp32(sim, ref)
    # single-pair run_tf('test', i, j) == the reference's per-pair loop (train.py:57-61)
    for i, j in ((0, 0), (1, 5), (4, 15)):
        s, secs = run_tf(data, dc, model, saver, sess, 'test', i, j)
        assert close_fp32(s, ref[i, j]) and secs > 0


@pytest.mark.parametrize('name', ['default', 'avg_ntn', 'dot'])
def test_training_losses_and_params_follow_reference_run(name):
    from graphembedding_b200.train import run_tf
    d, fl, data, dc, model, sess, saver = _setup(name)
    for it, st in enumerate(d['steps']):
        loss, _ = run_tf(data, dc, model, saver, sess, 'train', iter=it)
        assert abs(loss - st['loss']) <= 2e-4 * abs(st['loss']), (it, loss, st['loss'])
        vloss, _ = run_tf(data, dc, model, saver, sess, 'val', iter=it)
        assert abs(vloss - st['val_loss']) <= 1e-3 * abs(st['val_loss']), (it, vloss, st['val_loss'])
        got = model.engine.get_params().astype(np.float64)
        ref = np.array(st['params_after'])
        # Adam steps are ~lr per element; fp32 vs fp64 gradients flip only tiny-gradient elements
        assert np.max(np.abs(got - ref)) <= 0.2 * fl['learning_rate'] * (it + 1)
        assert np.mean(np.abs(got - ref)) <= 2e-3 * fl['learning_rate'] * (it + 1)


def test_unknown_tvt_raises_like_reference():
    from graphembedding_b200.train import run_tf
    d, fl, data, dc, model, sess, saver = _setup('default')
    with pytest.raises((RuntimeError, AssertionError)):
        run_tf(data, dc, model, saver, sess, 'bogus')


def test_tuning_sweep_runs_end_to_end(tmp_path):
    """two points of the sweep through the real train.main() (train + validate + all-pairs test +
    metrics on the device), CSV in the reference's layout."""
    from graphembedding_b200 import tuning
    grid = [(True, 0.6, 'sim_kernel', 0.01, 20, 0.25, 0.0), (False, 0.01, 'sigmoid', 0.01, 20, 0.25, 0.1)]
    path = str(tmp_path / 'sweep.csv')
    best_train, best_val = tuning.tune(file_name=path, grid=grid)
    lines = open(path).read().strip().split('\n')
    assert lines[0].split(',') == tuning.header and len(lines) == 1 + 2 + 5
    for ln in lines[1:3]:
        cells = ln.split(',')
        assert len(cells) == len(tuning.header)
        vals = [float(c) for c in cells[6:]]
        assert all(np.isfinite(vals)) and all(0.0 <= v <= 1.0 for v in vals[4:24])
    assert len(best_train) == len(best_val) == len(tuning.header) - 2
