"""world_size-2 gloo tests (CPU) of the multi-GPU host logic in graphembedding_b200/distributed.py:
database sharding + candidate all-gather + merge, and the data-parallel gradient all-reduce with the
global-batch target statistics.  The device kernels the GPU path plugs in (sgnn_topk /
sgnn_topk_merge / sgnn_train_step) are replaced by their host statements here; they are checked
against the same statements by the -m gpu tests."""
import os
import socket

import numpy as np
import pytest

torch = pytest.importorskip('torch')
import torch.distributed as dist  # noqa: E402
import torch.multiprocessing as mp  # noqa: E402

from util import O  # noqa: E402
from graphembedding_b200 import distributed as D  # noqa: E402


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _sim_matrix(m, n):
    rng = np.random.default_rng(17)
    sim = rng.integers(0, 20, size=(m, n)).astype(np.float32) / 19.0   # heavy ties across shards
    sim[2, :] = 0.25
    return sim


def _worker(rank, world, port, out):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        # ---- ranking: shard the database, local top-k, gather, merge ----
        m, n, k = 7, 1001, 10
        sim = _sim_matrix(m, n)
        lo, hi = D.shard_range(n, rank, world)
        local = sim[:, lo:hi]
        order = O.sort_id_mat_sim(local)[:, :k]                     # what sgnn_topk returns
        lv = torch.as_tensor(np.take_along_axis(local, order, 1).copy())
        li = torch.as_tensor(order.astype(np.int32).copy())
        cv, ci = D.gather_topk_candidates(lv, li, lo, dist)
        ids, vals = D.merge_topk_host(cv.numpy(), ci.numpy(), k)    # what sgnn_topk_merge returns
        ref = O.sort_id_mat_sim(sim)[:, :k]
        ok_rank = bool(np.array_equal(ids, ref)) and bool(np.array_equal(vals, np.take_along_axis(sim, ref, 1)))
        # ---- training: pair stream split, gradient all-reduce, global target statistics ----
        gb, npar = 12, 257
        rng = np.random.default_rng(5)
        per_pair_grad = rng.standard_normal((gb, npar)).astype(np.float32)   # stand-in for sgnn_train_step
        targets = rng.random(gb).astype(np.float32)
        mine = D.pair_slice(gb, rank, world)
        g = torch.as_tensor(per_pair_grad[mine].sum(axis=0))
        D.allreduce_gradients(g, dist)
        s1, s2 = D.global_target_stats(targets)
        ok_train = bool(np.allclose(g.numpy(), per_pair_grad.sum(axis=0), rtol=1e-5, atol=1e-5))
        ok_stats = abs(s1 - float(targets.astype(np.float64).sum())) < 1e-12
        out[rank] = (ok_rank, ok_train, ok_stats, [int(x) for x in mine])
    finally:
        dist.destroy_process_group()


def test_two_rank_ranking_merge_and_dp_allreduce():
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    assert set(out.keys()) == {0, 1}
    seen = []
    for r in range(world):
        ok_rank, ok_train, ok_stats, mine = out[r]
        assert ok_rank and ok_train and ok_stats, (r, out[r])
        seen += mine
    assert sorted(seen) == list(range(12))          # the pair stream is partitioned, nothing dropped


def test_shard_ranges_partition_the_database():
    for n in (0, 1, 7, 10000, 10001):
        for world in (1, 2, 3, 8):
            edges = [D.shard_range(n, r, world) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == n
            assert all(edges[r][1] == edges[r + 1][0] for r in range(world - 1))
            sizes = [hi - lo for lo, hi in edges]
            assert max(sizes) - min(sizes) <= 1


def test_merge_host_statement_is_reference_tie_order():
    rng = np.random.default_rng(3)
    sim = rng.integers(0, 5, size=(4, 64)).astype(np.float32)
    ids = np.broadcast_to(np.arange(64, dtype=np.int32), sim.shape)
    got, _ = D.merge_topk_host(sim, ids, 9)
    assert np.array_equal(got, O.sort_id_mat_sim(sim)[:, :9])


class _HostMetricEngine(object):
    """Host statements of sgnn_precision_at_k / sgnn_kendall_tau (the oracle's precision, scipy's tau-b: what the
    -m gpu tests hold the kernels to) behind the two engine methods sharded_full_row_metrics calls; the 2-GPU test
    (tests/multi_gpu_worker.py) plugs the kernels in."""

    def precision_at_ks(self, true_sim, pred_sim, ks):
        t, p = true_sim.numpy(), pred_sim.numpy()
        return torch.as_tensor(np.asarray(O.precision_at_ks(O.sort_id_mat_sim(t), t, O.sort_id_mat_sim(p), p, list(ks)), np.float64))

    def kendall_tau(self, true_sim, pred_sim):
        from scipy.stats import kendalltau
        tau = [kendalltau(a, b).statistic for a, b in zip(true_sim.numpy(), pred_sim.numpy())]
        return torch.as_tensor(np.asarray(tau, np.float64)), None


def _metric_worker(rank, world, port, out):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        m, n = 9, 203                                   # rows and columns that do not divide by the world size
        rng = np.random.default_rng(23)
        true = rng.random((m, n)).astype(np.float32)
        pred = (true + 0.3 * rng.standard_normal((m, n))).astype(np.float32)
        lo, hi = D.shard_range(n, rank, world)
        rows, (rlo, rhi) = D.redistribute_rows(torch.as_tensor(pred[:, lo:hi].copy()), dist)
        ok_rows = bool(np.array_equal(rows.numpy(), pred[rlo:rhi])) and (rlo, rhi) == D.shard_range(m, rank, world)
        eng = _HostMetricEngine()
        prec, tau = D.sharded_full_row_metrics(eng, torch.as_tensor(true[:, lo:hi].copy()), torch.as_tensor(pred[:, lo:hi].copy()),
                                               [1, 10, 50], dist)
        ref_p = eng.precision_at_ks(torch.as_tensor(true), torch.as_tensor(pred), [1, 10, 50]).numpy()
        ref_t, _ = eng.kendall_tau(torch.as_tensor(true), torch.as_tensor(pred))
        out[rank] = (ok_rows, bool(np.allclose(prec, ref_p, rtol=0, atol=1e-12)),
                     bool(np.allclose(tau, ref_t.numpy(), rtol=0, atol=1e-12)))
    finally:
        dist.destroy_process_group()


def test_two_rank_row_redistribution_and_full_row_metrics():
    """Kendall tau / precision@k of a database-sharded matrix = the single-process values after the
    all-to-all that gives every rank m / R whole rows."""
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_metric_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    for r in range(world):
        assert out[r] == (True, True, True), (r, out[r])
