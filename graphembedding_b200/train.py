"""Train / validate / test drivers with the reference's signatures
(model/Siamese/train.py: ``train_val`` :8-44, ``test`` :47-74, ``run_tf`` :77-102,
models_factory.py:5-11, main.py:14-27).

``run_tf`` keeps the per-call contract ``(last_output, seconds)``.  ``test`` replaces the
reference's m x n loop of single-pair ``sess.run`` calls (78 400 launches on AIDS700nef) by
one embedding pass over the m + n graphs, one all-pairs scoring call and on-device ranking;
the similarity goes to ``test_sim_mat[i][j]`` (the reference writes ``[i][i]``, train.py:68,
which discards all but one score per query -- deliberately not reproduced)."""
from time import time

import numpy as np

from .model_mse import SiameseGCNTNMSE, Token


class Session(object):
    """Stand-in for tf.Session: ``run(objs, feed_dict)`` dispatches to the model."""

    def __init__(self, model=None):
        self.model = model
        self.graph = None

    def run(self, objs, feed_dict=None):
        if not objs or not isinstance(objs[-1], Token):
            return []
        return self.model.run(objs, feed_dict)


class Saver(object):
    """No-op logging hooks (the reference's TensorBoard writer, saver.py:7-72, is out of scope)."""

    def __init__(self, sess=None):
        pass

    def proc_objs(self, objs, tvt):
        return objs

    def proc_outs(self, outs, tvt, iter):
        pass


def create_model(model, input_dim, **overrides):
    if model == 'siamese_gcntn_mse':
        return SiameseGCNTNMSE(input_dim, **overrides)
    if model == 'siamese_gcntn_hinge':
        raise NotImplementedError('siamese_gcntn_hinge is a stub in the reference (model_hinge.py:5-45)')
    raise RuntimeError('Unknown model {}'.format(model))


def run_tf(data, dist_calculator, model, saver, sess, tvt, test_id=None, train_id=None, iter=None):
    feed_dict = model.get_feed_dict(data, dist_calculator, tvt, test_id, train_id)
    if tvt == 'train':
        objs = [model.opt_op, model.loss]
    elif tvt == 'val':
        objs = [model.loss]
    elif tvt == 'test':
        objs = [model.pred_sim_without_act()]
    else:
        raise RuntimeError('Unknown train_val_test {}'.format(tvt))
    objs = saver.proc_objs(objs, tvt)
    t = time()
    outs = sess.run(objs, feed_dict=feed_dict)
    time_rtn = time() - t
    saver.proc_outs(outs, tvt, iter)
    if tvt == 'test':
        assert len(outs[-1]) == 1
        outs[-1] = model.apply_final_act_np(outs[-1][0])                    # numpy fp64 kernel, train.py:98-101
    return outs[-1], time_rtn


def train_val(data, dist_calculator, model, saver, sess, iters=None, early_stopping=None, verbose=True):
    from .config import FLAGS
    iters = FLAGS.iters if iters is None else iters
    early_stopping = FLAGS.early_stopping if early_stopping is None else early_stopping
    train_costs, train_times, val_costs, val_times = [], [], [], []
    for it in range(iters):
        c, s = run_tf(data, dist_calculator, model, saver, sess, 'train', iter=it)
        train_costs.append(c)
        train_times.append(s)
        vc, vs = run_tf(data, dist_calculator, model, saver, sess, 'val', iter=it)
        val_costs.append(vc)
        val_times.append(vs)
        if verbose:
            print('Iter:', '%04d' % (it + 1), 'train_loss=', '{:.5f}'.format(c),
                  'time=', '{:.2f}msec'.format(s * 1000), 'val_loss=', '{:.5f}'.format(vc),
                  'time=', '{:.2f}msec'.format(vs * 1000))
        if early_stopping and it > early_stopping and \
                val_costs[-1] > np.mean(val_costs[-(early_stopping + 1):-1]):
            print('Early stopping...')
            break
    print('Optimization Finished!')
    return train_costs, train_times, val_costs, val_times


def all_pairs_similarity(data, model, k=10):
    """Similarity of every (test graph i, original train graph j): one embedding pass, one
    all-pairs call.  Returns (sim_mat float64 [m,n], topk ids, seconds)."""
    import torch
    m, n = data.m_n()
    eng = model.engine
    _, dev, offsets, _ = model._ensure_resident(data)
    q_ids = np.array([model._gid(offsets, data.test_data.get_graph(i)) for i in range(m)], np.int64)
    d_ids = np.array([model._gid(offsets, data.get_orig_train_graph(j)) for j in range(n)], np.int64)
    eng.reserve(max_queries=m, max_db=n)
    if eng.has_head:   # SimGNN-style head: the histogram needs each side's node rows next to the embeddings
        host = model._ensure_resident(data)[3]
        qb, db = host.select(q_ids).to(eng.device), host.select(d_ids).to(eng.device)
        torch.cuda.synchronize()
        t = time()
        emb_q, rows_q = eng.embed_nodes(qb)
        emb_d, rows_d = eng.embed_nodes(db)
        sim = eng.score_all_pairs_nodes(emb_q, rows_q, qb, emb_d, rows_d, db)
    else:
        torch.cuda.synchronize()
        t = time()
        emb = eng.embed(dev)
        emb_q = emb[torch.as_tensor(q_ids, device=eng.device)].contiguous()
        emb_d = emb[torch.as_tensor(d_ids, device=eng.device)].contiguous()
        sim = eng.score_all_pairs(emb_q, emb_d)
    ids, _ = eng.topk(sim, min(k, n - 1)) if n > 1 else (None, None)
    torch.cuda.synchronize()
    secs = time() - t
    return sim.cpu().numpy().astype(np.float64), (None if ids is None else ids.cpu().numpy()), secs


def test(data, dist_calculator, model, saver, sess, eval_obj=None):
    m, n = data.m_n()
    run_tf(data, dist_calculator, model, saver, sess, 'test', 0, 0)       # flush the pipeline (train.py:54-55)
    test_sim_mat, _, secs = all_pairs_similarity(data, model)
    test_time_mat = np.full((m, n), secs * 1000.0 / (m * n))             # msec per pair
    if eval_obj is None:
        from .eval import Eval
        eval_obj = Eval.from_data(data, dist_calculator, model)
    print('Evaluating...')
    from . import metrics
    metrics.use_engine(model.engine)          # precision@k / Kendall tau on the device
    results = eval_obj.eval_test(model.__class__.__name__.lower().replace('gcntnmse', '_gcntn_mse'),
                                 test_sim_mat, test_time_mat)
    metrics.use_engine(None)
    print('Results generated with {} metrics'.format(len(results)))
    return results


def main(**overrides):
    """main.py:14-27."""
    from .config import FLAGS, ensure_parsed
    from .data_siamese import SiameseModelData
    from .dist_calculator import DistCalculator
    from .utils import check_flags
    ensure_parsed()
    check_flags()
    data = SiameseModelData()
    dist_calculator = DistCalculator(FLAGS.dataset, FLAGS.dist_metric, FLAGS.dist_algo)
    model = create_model(FLAGS.model, data.input_dim(), **overrides)
    sess = Session(model)
    saver = Saver(sess)
    train_costs, train_times, val_costs, val_times = train_val(data, dist_calculator, model, saver, sess)
    results = test(data, dist_calculator, model, saver, sess)
    return train_costs, train_times, val_costs, val_times, results
