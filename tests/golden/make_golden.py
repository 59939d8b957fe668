#!/usr/bin/env python3
"""Generate the golden fixtures under tests/golden/ by RUNNING THE REFERENCE in this
container (it cannot travel to the GPU box; the fixtures do).

    python tests/golden/make_golden.py            # all parts

Parts
  host   : the reference's own host modules, imported unmodified from /root/reference:
           model/Siamese/graphs.py (ModelGraph, NodeFeatureOneHotEncoder), samplers.py,
           src/similarity.py, src/distance.py (normalized_dist), src/results.py + metrics.py.
  model  : the reference's own model code (config.py, inits.py, layers.py, layers_factory.py,
           models.py, model_mse.py, data_siamese.py, dist_calculator.py), unmodified, executed
           on oracle/tf1_shim.py (a torch-backed stand-in for the TF-1 API; TensorFlow itself is
           not installed).  One subprocess per model configuration because flags and layer
           names are process-global in the reference.

Synthetic inputs come from graphembedding_b200.data (the real datasets are unavailable).
"""
import json
import os
import subprocess
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = '/root/reference'
sys.path.insert(0, ROOT)


def _ref_paths():
    sys.path.insert(0, REF + '/src')
    sys.path.insert(0, REF + '/model/Siamese')
    sys.modules.setdefault('bs4', types.ModuleType('bs4'))
    sys.modules['bs4'].BeautifulSoup = object


def _graphs_to_lists(gs):
    out = []
    for g in gs:
        nodes = list(g.nodes())
        pos = {n: i for i, n in enumerate(nodes)}
        out.append(dict(gid=int(g.graph['gid']), types=[g.nodes[n]['type'] for n in nodes],
                        edges=[[pos[u], pos[v]] for u, v in g.edges()]))
    return out


# ---------------------------------------------------------------------------------------
def part_host():
    _ref_paths()
    import networkx as nx
    from graphs import ModelGraph, NodeFeatureOneHotEncoder          # reference
    from samplers import RandomSampler, DistributionSampler           # reference
    from similarity import create_sim_kernel                          # reference
    from distance import normalized_dist                              # reference
    import results as R                                               # reference
    import metrics as MT                                              # reference
    from graphembedding_b200 import data as D

    # --- pre-processing -----------------------------------------------------------------
    gs = D.synthetic_corpus(40, 1, 10, seed=501) + D.synthetic_corpus(8, 11, 30, seed=502, gid_base=100)
    iso = nx.Graph(); iso.add_node('0', type='C'); iso.add_node('1', type='O'); iso.graph['gid'] = 999
    gs.append(iso)  # two isolated nodes: degree-0 rows
    enc = NodeFeatureOneHotEncoder(gs, 'type')
    rec = dict(feat_idx_dic=enc.feat_idx_dic, input_dim=int(enc.input_dim()), graphs=_graphs_to_lists(gs), ref=[])
    for g in gs:
        mg = ModelGraph(g, enc)
        fc, fv, fs = mg.get_node_inputs()
        lc, lv, ls = mg.get_laplacians()[0]
        rec['ref'].append(dict(feat_coords=fc.tolist(), feat_vals=fv.tolist(), feat_shape=list(fs),
                               nnz_shape=list(mg.get_node_inputs_num_nonzero()),
                               lap_coords=lc.tolist(), lap_vals_hex=[float(x).hex() for x in lv],
                               lap_shape=list(ls)))
    json.dump(rec, open(HERE + '/preprocess.json', 'w'))

    # --- samplers -----------------------------------------------------------------------
    class G(object):  # minimal stand-in for ModelGraph inside the sampler
        def __init__(self, i, nxg):
            self.i, self.nxgraph = i, nxg
    srec = {}
    for L in (3, 5, 52):
        items = [G(i, gs[i % len(gs)]) for i in range(L)]
        s = RandomSampler(items, -1, False)
        srec['random_%d' % L] = [[a.i, b.i] for a, b in (s.get_pair() for _ in range(4 * L + 7))]
    for L, sn in ((40, -1), (40, 3)):
        items = [G(i, gs[i]) for i in range(L)]
        s = DistributionSampler(items, sn, False)
        srec['density_%d_%d' % (L, sn)] = [[a.i, b.i] for a, b in (s.get_pair() for _ in range(30))]
    json.dump(srec, open(HERE + '/samplers.json', 'w'))

    # --- similarity / distance ----------------------------------------------------------
    ds = np.linspace(0, 4, 41)
    krec = {'d': ds.tolist()}
    for y in (0.6, 0.3, 0.2, 0.01):
        krec['gauss_%g' % y] = [float(v).hex() for v in create_sim_kernel('gaussian', y).dist_to_sim_np(ds)]
    krec['identity'] = create_sim_kernel('identity').dist_to_sim_np(ds).tolist()
    krec['normalized_dist'] = [[d, gs[i].number_of_nodes(), gs[j].number_of_nodes(),
                                float(normalized_dist(d, gs[i], gs[j])).hex()]
                               for d, i, j in ((3, 0, 1), (7, 2, 5), (0, 3, 3), (11, 41, 44))]
    json.dump(krec, open(HERE + '/similarity.json', 'w'))

    # --- ranking + metrics (results.py / metrics.py) ---------------------------------------
    rng = np.random.default_rng(77)
    m, n = 9, 31
    qs = D.synthetic_corpus(m, 2, 10, seed=601)
    dbs = D.synthetic_corpus(n, 2, 10, seed=602)
    dist = rng.integers(0, 9, size=(m, n)).astype(np.float64)          # integer GEDs: many ties
    dist[4, :] = 3.0                                                    # fully tied query
    norm = np.array([[normalized_dist(dist[i, j], qs[i], dbs[j]) for j in range(n)] for i in range(m)])
    pred = np.round(rng.uniform(0, 1, size=(m, n)), 2)                  # ties in predictions too
    pred[2, :] = 0.25

    class TrueR(R.PairwiseGEDModelResult):
        def __init__(self):
            self.model_ = 'astar'; self.dataset = 'syn'
            self.dist_mat_ = dist; self.dist_norm_mat_ = norm
            self.time_mat_ = np.zeros((m, n))
            self.sort_id_mat_ = np.argsort(self.dist_mat_, kind='mergesort')          # results.py:143
            self.dist_norm_sort_id_mat_ = np.argsort(self.dist_norm_mat_, kind='mergesort')
    tr = TrueR()
    pr = R.SiameseModelResult('syn', 'siamese_gcntn_mse', sim_mat=pred, time_mat=np.ones((m, n)))
    ks = [1, 2, 3, 5, 10, 20, 30]
    mrec = dict(m=m, n=n, dist=dist.tolist(), norm=norm.tolist(), pred=pred.tolist(), ks=ks,
                q_nodes=[g.number_of_nodes() for g in qs], db_nodes=[g.number_of_nodes() for g in dbs],
                pred_sort_id=pr.sort_id_mat().tolist())
    for nm in (True, False):
        key = 'norm' if nm else 'nonorm'
        mrec['true_sort_id_' + key] = tr.sort_id_mat(nm).tolist()
        mrec['prec_' + key] = MT.precision_at_ks(tr, pr, nm, ks).tolist()
        mrec['prec_all_' + key] = MT.precision_at_ks(tr, pr, nm, list(range(1, n))).tolist()
        mrec['mrr_' + key] = float(MT.mean_reciprocal_rank(tr, pr, nm))
        mrec['mse_' + key] = float(MT.mean_squared_error(tr, pr, 'gaussian', 0.6, nm))
        mrec['true_topk_incl_' + key] = [[tr.top_k_ids(i, k, nm, True).tolist() for k in ks] for i in range(m)]
        mrec['true_rank_' + key] = [[int(tr.ranking(i, j, nm)) for j in range(n)] for i in range(m)]
    mrec['pred_topk'] = [[pr.top_k_ids(i, k, False, False).tolist() for k in ks] for i in range(m)]
    mrec['pred_rank'] = [[int(pr.ranking(i, j, False)) for j in range(n)] for i in range(m)]
    mrec['time'] = float(MT.average_time(pr))
    json.dump(mrec, open(HERE + '/metrics.json', 'w'))
    print('host goldens written')


# ---------------------------------------------------------------------------------------
MODEL_CONFIGS = {
    # the reference's default flags (config.py:44-66), dropout forced to 0 (F8)
    'default': ['--batch_size=4'],
    # the tuning.py:74-93 architecture
    'avg_ntn': ['--batch_size=3', '--num_layers=4', '--layer_2=Average',
                '--layer_3=NTN:input_dim=16,feature_map_dim=10,inneract=relu,dropout=True,bias=True'],
    # Dot pair layer, identity final act, unnormalised distance
    'dot': ['--batch_size=5', '--num_layers=4', '--layer_2=Average', '--layer_3=Dot',
            '--final_act=sigmoid', '--dist_norm=False', '--yeta=0.05'],
}


def part_model(name):
    os.environ.setdefault('TF1_SHIM_DTYPE', 'float64')
    from oracle import tf1_shim
    tf1_shim.install()
    _ref_paths()
    import torch
    torch.manual_seed(0)
    from graphembedding_b200 import data as D
    import utils as ref_utils                       # reference src/utils.py
    import data as ref_data                         # reference src/data.py
    ref_data.load = lambda f: None                  # no pickle cache in the read-only tree
    ref_data.save = lambda f, o: None

    train_gs = D.synthetic_corpus(16, 2, 10, seed=701)
    test_gs = D.synthetic_corpus(5, 2, 10, seed=702, gid_base=16)

    class Fake(object):
        def __init__(self, graphs):
            self.graphs = graphs
    import config                                    # reference config.py (flags on absl)
    FLAGS = config.FLAGS
    FLAGS(['make_golden', '--dataset=aids80nef', '--dropout=0', '--log=False', '--iters=3'] + MODEL_CONFIGS[name])
    import data_siamese                              # reference
    data_siamese.load_data = lambda ds, train: Fake(list(train_gs) if train else list(test_gs))
    import dist_calculator as ref_dc                 # reference
    ref_dc.load = lambda f: None
    ref_dc.save = lambda f, o: None
    from graphembedding_b200.dist_calculator import proxy_ged
    ref_dc.ged = lambda g1, g2, algo: proxy_ged(g1, g2)
    from utils_siamese import check_flags
    from models_factory import create_model          # reference (imports model_mse, model_hinge)

    check_flags()
    data = data_siamese.SiameseModelData()
    dc = ref_dc.DistCalculator(FLAGS.dataset, FLAGS.dist_metric, FLAGS.dist_algo)
    model = create_model(FLAGS.model, data.input_dim())
    sess = tf1_shim.Session()
    sess.run(tf1_shim.global_variables_initializer())

    # deterministic weights: glorot ranges as inits.py, numpy RNG; small non-zero biases
    rng = np.random.default_rng(2024)
    flat = []
    var_names = []
    for layer in model.layers:
        for vname, var in layer.vars.items():
            shape = tuple(var.value.shape)
            if vname == 'bias':
                val = rng.uniform(-0.05, 0.05, size=shape)
            else:
                r = np.sqrt(6.0 / (shape[0] + shape[1]))
                val = rng.uniform(-r, r, size=shape)
            val = val.astype(np.float32).astype(np.float64)
            var.assign_value(val)
            flat.append(val.ravel()); var_names.append([layer.get_name(), vname, list(shape)])
    flat0 = np.concatenate(flat)

    def cur_params():
        return np.concatenate([v.numpy().ravel() for layer in model.layers for v in layer.vars.values()])

    # ---- test(): m x n single-pair runs (train.py:47-74, 77-102) ----
    m, n = data.m_n()
    raw = np.zeros((m, n)); sim = np.zeros((m, n))
    for i in range(m):
        for j in range(n):
            fd = model.get_feed_dict(data, dc, 'test', i, j)
            outs = sess.run([model.pred_sim_without_act()], feed_dict=fd)
            assert len(outs[-1]) == 1
            raw[i, j] = outs[-1][0]
            sim[i, j] = model.apply_final_act_np(outs[-1][0])
    db_gids = [data.get_orig_train_graph(j).get_nxgraph().graph['gid'] for j in range(n)]
    q_gids = [data.test_data.get_graph(i).get_nxgraph().graph['gid'] for i in range(m)]

    # ---- train_val(): FLAGS.iters steps (train.py:8-44) with the sampler calls logged ----
    log = []
    orig_pair = data.get_graph_pair

    def logged(tvt):
        g1, g2 = orig_pair(tvt)
        log.append([tvt, g1.get_nxgraph().graph['gid'], g2.get_nxgraph().graph['gid']])
        return g1, g2
    data.get_graph_pair = logged
    steps = []
    for it in range(FLAGS.iters):
        del log[:]
        fd = model.get_feed_dict(data, dc, 'train', None, None)
        calls = [list(c) for c in log]
        dists = np.asarray(fd[model.dists]).ravel().tolist()
        norm_dists = np.asarray(fd[model.norm_dists]).ravel().tolist()
        before = cur_params()
        outs = sess.run([model.opt_op, model.loss], feed_dict=fd)
        grads = np.concatenate([model.optimizer.last_grads[v.name].numpy().ravel()
                                for layer in model.layers for v in layer.vars.values()])
        del log[:]
        fdv = model.get_feed_dict(data, dc, 'val', None, None)
        vcalls = [list(c) for c in log]
        vloss = sess.run([model.loss], feed_dict=fdv)[-1]
        steps.append(dict(sampler_calls=calls, dists=dists, norm_dists=norm_dists,
                          params_before=before.tolist(), loss=float(outs[-1]), grads=grads.tolist(),
                          params_after=cur_params().tolist(), val_sampler_calls=vcalls,
                          val_dists=np.asarray(fdv[model.dists]).ravel().tolist(),
                          val_norm_dists=np.asarray(fdv[model.norm_dists]).ravel().tolist(),
                          val_loss=float(vloss)))
    flags = {k: FLAGS.flag_values_dict()[k] for k in
             ['batch_size', 'num_layers', 'yeta', 'final_act', 'dist_norm', 'weight_decay', 'learning_rate',
              'sim_kernel', 'valid_percentage'] + ['layer_%d' % i for i in range(FLAGS.num_layers)]}
    rec = dict(flags=flags, feat_idx_dic=data.node_feat_encoder.feat_idx_dic, input_dim=int(data.input_dim()),
               train_graphs=_graphs_to_lists(train_gs), test_graphs=_graphs_to_lists(test_gs),
               var_names=var_names, params0=flat0.tolist(), test_raw=raw.tolist(), test_sim=sim.tolist(),
               q_gids=q_gids, db_gids=db_gids, steps=steps, shim_dtype=os.environ['TF1_SHIM_DTYPE'])
    json.dump(rec, open(HERE + '/model_%s.json' % name, 'w'))
    print('model golden', name, 'written: losses', [s['loss'] for s in steps])


# ---------------------------------------------------------------------------------------
def part_attention():
    """The reference's Attention layer (layers.py:143-160) is unreachable through its factory, but the
    class itself runs: layers.Attention(input_dim=d)._call_one_mat on the torch-backed TF-1 stand-in, fp64."""
    os.environ.setdefault('TF1_SHIM_DTYPE', 'float64')
    from oracle import tf1_shim
    tf1_shim.install()
    _ref_paths()
    import config                                    # reference flags (layers.py reads FLAGS.dropout)
    config.FLAGS(['make_golden', '--dropout=0', '--log=False'])
    import layers as ref_layers                      # reference
    import tensorflow as tf                          # the shim
    rng = np.random.default_rng(77)
    rec = []
    for d, n in ((16, 7), (16, 1), (12, 10), (8, 30)):
        layer = ref_layers.Attention(input_dim=d)
        x_ph = tf.placeholder(tf.float32)
        out_node = layer._call_one_mat(x_ph)         # the reference's graph for one N x D matrix
        sess = tf1_shim.Session()
        sess.run(tf1_shim.global_variables_initializer())
        W = rng.uniform(-0.6, 0.6, size=(d, d)).astype(np.float32).astype(np.float64)
        layer.vars['weights'].assign_value(W)
        x = rng.standard_normal((n, d)).astype(np.float32).astype(np.float64)
        out = np.asarray(sess.run(out_node, feed_dict={x_ph: x}), np.float64).reshape(-1)
        rec.append(dict(d=d, n=n, weights=W.tolist(), x=x.tolist(), out=out.tolist()))
    json.dump(rec, open(HERE + '/attention.json', 'w'))
    print('attention golden written:', [len(r['out']) for r in rec])


if __name__ == '__main__':
    if len(sys.argv) > 1 and sys.argv[1] == 'host':
        part_host()
    elif len(sys.argv) > 2 and sys.argv[1] == 'model':
        part_model(sys.argv[2])
    elif len(sys.argv) > 1 and sys.argv[1] == 'attention':
        part_attention()
    else:
        subprocess.check_call([sys.executable, __file__, 'host'])
        for nm in MODEL_CONFIGS:
            subprocess.check_call([sys.executable, __file__, 'model', nm])
        subprocess.check_call([sys.executable, __file__, 'attention'])
