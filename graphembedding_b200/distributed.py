"""Multi-GPU plumbing of the two places the path shards; one process per GPU,
``torch.distributed`` (NCCL on the GPUs, gloo in the CPU tests) for the collectives.

* Training is data-parallel over the pair stream: pair ``t`` of the global mini-batch goes to rank
  ``t mod R`` (so R = 1 and R = 8 see the same global batch), every rank runs ``sgnn_train_step`` on
  its slice, ONE all-reduce sums the flat gradient blob (a single bucket: 11-40 KB), and every rank
  applies the identical ``sgnn_apply_adam``.  The reference-mode loss (model_mse.py:145-151, SURVEY
  F6) needs the target sum / square sum of the GLOBAL batch: targets are host data, so they are
  computed on the host from the global batch -- no extra collective.
* All-pairs ranking shards the DATABASE contiguously; queries are replicated.  Each rank scores its
  m x n_r tile and keeps a local top-k; the (score, GLOBAL id) candidates of all ranks are merged
  with the reference's tie order (results.py:221-224: equal scores -> higher index first).  On the
  GPUs of one node the exchange is ``PeerTopk``: every rank pushes its candidates into every peer's
  buffer over NVLink peer memory and merges on arrival (two kernels, sgnn_peer.cuh).  The
  all-gather + ``sgnn_topk_merge`` path is the fallback (no peer access) and what the gloo tests run.

Nothing here touches the reference's semantics; the functions are written against plain tensors so
that the world_size-2 gloo tests (tests/test_dist_gloo.py) exercise exactly the code the GPU path
runs."""
import numpy as np


def shard_range(n, rank, world):
    """Contiguous database shard [lo, hi) of rank ``rank``: sizes differ by at most one."""
    base, rem = divmod(int(n), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def pair_slice(global_batch, rank, world):
    """Indices of the global mini-batch that rank ``rank`` trains on (t mod R == rank)."""
    return np.arange(rank, int(global_batch), int(world))


def global_target_stats(target_sim_global):
    """(sum, sum of squares) of the global batch's target similarities in float64, the two
    scalars ``sgnn_train_step`` needs for the reference-mode loss on a shard of the batch."""
    t = np.asarray(target_sim_global, np.float64)
    return float(t.sum()), float((t * t).sum())


def allreduce_gradients(grads, dist=None, group=None):
    """Sum the flat gradient blob over ranks in place (one bucket)."""
    if dist is not None and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(grads, op=dist.ReduceOp.SUM, group=group)
    return grads


def gather_topk_candidates(local_vals, local_ids, id_offset, dist=None, group=None):
    """All-gather the per-shard top-k and lay the candidates out as [m, world * k] with GLOBAL
    database ids (local id + the shard's offset).  Returns (cand_vals, cand_ids)."""
    import torch
    ids = local_ids + int(id_offset)
    world = dist.get_world_size(group) if dist is not None and dist.is_initialized() else 1
    if world == 1:
        return local_vals, ids
    m, k = local_vals.shape
    # ONE collective: scores (bit pattern) and ids travel together as int32 [2, m, k]
    packed = torch.stack([local_vals.contiguous().view(torch.int32), ids.to(torch.int32)], dim=0)
    got = torch.empty((world * 2 * m, k), dtype=torch.int32, device=packed.device)   # rank-major concat
    dist.all_gather_into_tensor(got, packed.view(2 * m, k), group=group)
    got = got.view(world, 2, m, k)
    cv = got[:, 0].permute(1, 0, 2).reshape(m, world * k).contiguous().view(local_vals.dtype)
    ci = got[:, 1].permute(1, 0, 2).reshape(m, world * k).contiguous()
    return cv, ci


class PeerTopk(object):
    """Peer-memory candidate exchange of one process group (one process per GPU, one node).

    Setup is a host-side all-gather of 64-byte CUDA IPC handles; each step is
    ``engine.topk_exchange`` (a push kernel + a merge kernel, no library collective).  Every rank
    must call ``exchange`` the same number of times."""

    def __init__(self, engine, m_cap, k_cap, dist, group=None):
        self.engine, self.dist, self.group = engine, dist, group
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        mine = engine.peer_create(world, rank, m_cap, k_cap)
        handles = [None] * world
        dist.all_gather_object(handles, mine, group=group)
        engine.peer_connect(handles)
        dist.barrier(group=group)

    def exchange(self, local_vals, local_ids, id_offset, out_ids=None, out_vals=None):
        return self.engine.topk_exchange(local_vals, local_ids, id_offset, out_ids=out_ids, out_vals=out_vals)

    def barrier(self):
        """Device-side, stream-ordered barrier between the ranks (no host synchronisation)."""
        self.engine.peer_barrier()

    def adam_allreduce(self, grad_scale=1.0, loss=None):
        """DP optimiser step: gradient push + Adam on the rank-ordered sum of the world blobs."""
        self.engine.apply_adam_allreduce(grad_scale, loss)

    def close(self):
        import torch
        torch.cuda.synchronize()
        self.dist.barrier(group=self.group)   # no peer is still writing into / reading from a buffer
        self.engine.peer_destroy()


def merge_topk_host(cand_vals, cand_ids, k):
    """Host statement of the merge ``sgnn_topk_merge`` performs on device: descending score, ties
    -> higher GLOBAL id first (a stable ascending mergesort on (score, id) reversed,
    results.py:221-224).  Used by the CPU tests; the GPU path calls the kernel."""
    v = np.asarray(cand_vals)
    i = np.asarray(cand_ids)
    order = np.lexsort((i, v), axis=1)[:, ::-1][:, :k]
    return np.take_along_axis(i, order, 1), np.take_along_axis(v, order, 1)


def sharded_rank(engine, queries_dev, db_shard_dev, id_offset, k, dist=None, group=None, emb_q=None,
                 emb_db=None, sim=None, peer=None, ids=None, vals=None, out_ids=None, out_vals=None):
    """One database-sharded ranking step on this rank: embed, score the local tile, local top-k,
    candidate exchange + merge (``peer``: a PeerTopk; else all-gather + merge kernel).  Returns
    (ids, vals) of the global top-k (identical on every rank)."""
    if peer is not None:   # one replayable unit: local ranking + push + merge
        return engine.rank_sharded(queries_dev, db_shard_dev, k, id_offset, emb_q=emb_q, emb_db=emb_db, sim=sim,
                                   ids=ids, vals=vals, out_ids=out_ids, out_vals=out_vals)
    ids, vals, sim = engine.rank_device(queries_dev, db_shard_dev, k, emb_q=emb_q, emb_db=emb_db, sim=sim,
                                        ids=ids, vals=vals)
    world = dist.get_world_size(group) if dist is not None and dist.is_initialized() else 1
    if world == 1:
        return (ids, vals) if id_offset == 0 else (ids + int(id_offset), vals)
    cv, ci = gather_topk_candidates(vals, ids, id_offset, dist, group)
    return engine.topk_merge(cv, ci, k)


def data_parallel_step(engine, batch_dev, g1, g2, target_sim, global_batch, target_sum, target_sqsum, loss,
                       dist=None, group=None, grads=None, peer=None):
    """One DP optimiser step: local fwd+bwd, gradient all-reduce, identical Adam on every rank.
    ``peer`` (a PeerTopk of this engine): the all-reduce is the peer-memory push fused into Adam.
    The dropout masks are keyed on the pair's index in the GLOBAL batch (rank r owns the pairs t = r (mod R),
    ``pair_slice``): with the same seed on every rank, R = 1 and R = 8 draw the same masks for the same global batch."""
    if dist is not None and dist.is_initialized():
        engine.set_pair_offset(dist.get_rank(group), dist.get_world_size(group))
    engine.train_step(batch_dev, g1, g2, target_sim, global_batch, target_sum, target_sqsum, loss=loss)
    if peer is not None:
        peer.adam_allreduce(1.0, loss)
        return loss
    if grads is None:
        grads = engine.device_buffers()[1]
    allreduce_gradients(grads, dist, group)
    engine.apply_adam(1.0, loss)
    return loss


def redistribute_rows(local_cols, dist=None, group=None):
    """Database-sharded [m, n_r] column blocks -> this rank's full rows [m_r, n] (the ranking metrics
    that need WHOLE rows -- Kendall tau, precision@k beyond the merged top-k, MRR -- run on m / R rows per rank).
    Rank r receives rows ``shard_range(m, r, R)`` of every rank's block (one all-to-all; column order = rank order,
    i.e. the global database order of a ``shard_range`` split).  Returns (rows [m_r, n], (row_lo, row_hi))."""
    import torch
    world = dist.get_world_size(group) if dist is not None and dist.is_initialized() else 1
    m = int(local_cols.shape[0])
    if world == 1:
        return local_cols.contiguous(), (0, m)
    rank = dist.get_rank(group)
    n_local = torch.tensor([int(local_cols.shape[1])], dtype=torch.int64, device=local_cols.device)
    widths = [torch.zeros_like(n_local) for _ in range(world)]
    dist.all_gather(widths, n_local, group=group)
    widths = [int(w.item()) for w in widths]
    lo, hi = shard_range(m, rank, world)
    send = [local_cols[slice(*shard_range(m, r, world))].contiguous() for r in range(world)]
    recv = [torch.empty((hi - lo, widths[r]), dtype=local_cols.dtype, device=local_cols.device) for r in range(world)]
    if dist.get_backend(group) == 'nccl':
        dist.all_to_all(recv, send, group=group)
    else:   # gloo has no all_to_all: point-to-point (the CPU tests)
        ranks = list(range(world)) if group is None else dist.get_process_group_ranks(group)
        reqs = []
        for r in range(world):
            if r == rank:
                recv[r].copy_(send[r])
            else:
                reqs.append(dist.isend(send[r], dst=ranks[r], group=group))
                reqs.append(dist.irecv(recv[r], src=ranks[r], group=group))
        for q in reqs:
            q.wait()
    return torch.cat(recv, dim=1), (lo, hi)


def sharded_full_row_metrics(engine, true_local, pred_local, ks, dist=None, group=None):
    """precision@k (all k in ks) and Kendall tau-b of a database-sharded similarity matrix: the [m, n_r] blocks of the
    true and the predicted similarities are redistributed to m / R full rows per rank (``redistribute_rows``), the
    metric kernels (sgnn_precision_at_k, sgnn_kendall_tau) run on the local rows, and the per-row results are
    combined: tau gathered in row order, precision averaged with the row counts as weights (metrics.py:14-35 takes
    the mean over queries).  Returns (precision [len(ks)] float64 numpy, tau [m] float64 numpy), identical on every rank."""
    import torch
    world = dist.get_world_size(group) if dist is not None and dist.is_initialized() else 1
    t_rows, (lo, hi) = redistribute_rows(true_local, dist, group)
    p_rows, _ = redistribute_rows(pred_local, dist, group)
    m = int(true_local.shape[0])
    if hi > lo:
        prec = engine.precision_at_ks(t_rows, p_rows, ks).to(torch.float64) * float(hi - lo)
        tau, _ = engine.kendall_tau(t_rows, p_rows)
    else:
        prec = torch.zeros(len(ks), dtype=torch.float64, device=true_local.device)
        tau = torch.zeros(0, dtype=torch.float64, device=true_local.device)
    if world == 1:
        return (prec / m).cpu().numpy(), tau.cpu().numpy()
    dist.all_reduce(prec, group=group)
    most = (m + world - 1) // world          # shard_range gives the first ranks one row more: pad to a common length
    pad = torch.full((most,), float('nan'), dtype=torch.float64, device=tau.device)
    pad[: hi - lo] = tau
    parts = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad, group=group)
    taus = torch.cat([parts[r][: shard_range(m, r, world)[1] - shard_range(m, r, world)[0]] for r in range(world)])
    return (prec / m).cpu().numpy(), taus.cpu().numpy()
