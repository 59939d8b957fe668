"""Ranking / regression metrics with the reference's signatures (``src/metrics.py``:
``precision_at_ks`` :14-35, ``mean_reciprocal_rank`` :38-64, ``mean_squared_error`` :67-82,
``average_time`` :85-91) plus ``kendall_tau`` (absent in the reference; defined as
scipy's tau-b on each query row).

The reference evaluates O(m n^2) Python set intersections when ks = 1..n-1
(``src/exp.py:337-357``); here precision@k for ALL k of one query is one pass: an item
counts towards p@k once k >= max(its predicted position, its tie-inclusive true position)."""
import numpy as np

_ENGINE = None   # SgnnEngine whose device kernels evaluate precision@k / Kendall tau (set by use_engine)
DEVICE_MAX_N = 10000   # row length the shared-memory metric kernels accept (sgnn_rank.cuh)


def use_engine(engine):
    """Route precision_at_ks / kendall_tau through the device kernels of ``engine``
    (sgnn_precision_at_k / sgnn_kendall_tau); ``None`` restores the host (numpy) mirror of the
    reference's own host implementation."""
    global _ENGINE
    _ENGINE = engine


def _score_rows_f32(r, norm):
    """Row scores where higher = more similar, as float32 (distances are negated).  Ties and
    strict orderings of the float64 rows survive the cast: distances are small-integer ratios
    2d/(n1+n2) and similarities come from the float32 device path."""
    s = np.asarray(r.dist_sim_mat(norm), np.float64)
    if hasattr(r, 'dist_mat'):
        s = -s
    return np.ascontiguousarray(s, np.float32)


_HOST_NOTED = False


def _on_device(true_r, pred_r, norm):
    global _HOST_NOTED
    if _ENGINE is None:
        # Explicit, not silent: these are the reference's HOST functions (src/metrics.py is numpy / Python), so a
        # caller that never registered an engine gets the host mirror -- and is told so once.  The training /
        # evaluation entry points (train.main, tuning) register the engine; the product's scoring path never comes here.
        if not _HOST_NOTED:
            _HOST_NOTED = True
            import warnings
            warnings.warn('graphembedding_b200.metrics: no engine registered (metrics.use_engine(engine)): '
                          'precision@k / Kendall tau are evaluated by the host mirror of src/metrics.py, not by the '
                          'device kernels', stacklevel=3)
        return None
    m, n = true_r.m_n()
    if n > DEVICE_MAX_N or n < 2:
        return None
    import torch
    t = torch.as_tensor(_score_rows_f32(true_r, norm), device=_ENGINE.device)
    p = torch.as_tensor(_score_rows_f32(pred_r, norm), device=_ENGINE.device)
    return t, p


class Metric(object):
    def __init__(self, name, ylabel):
        self.name = name
        self.ylabel = ylabel

    def __str__(self):
        return self.name


def _tie_inclusive_position(sort_ids, scores):
    """For every item j of a row: the smallest k such that j is inside the tie-inclusive
    top-k (results.py:73-82) = 1 + index of the first element of j's tie run."""
    m, n = sort_ids.shape
    sc = np.take_along_axis(scores, sort_ids, 1)
    new_run = np.ones((m, n), bool)
    new_run[:, 1:] = sc[:, 1:] != sc[:, :-1]
    run_start = np.maximum.accumulate(np.where(new_run, np.arange(n)[None, :], 0), axis=1)
    pos = np.empty((m, n), np.int64)
    np.put_along_axis(pos, sort_ids, run_start + 1, 1)
    return pos


def precision_at_ks(true_r, pred_r, norm, ks, print_ids=[]):
    m, n = true_r.m_n()
    assert true_r.m_n() == pred_r.m_n()
    ks = list(ks)
    for k in ks:
        assert type(k) is int and 0 < k < n
    dev = _on_device(true_r, pred_r, norm)
    if dev is not None and not print_ids:
        return _ENGINE.precision_at_ks(dev[0], dev[1], ks).cpu().numpy()
    t_pos = _tie_inclusive_position(true_r.sort_id_mat(norm), true_r.dist_sim_mat(norm))
    p_ids = pred_r.sort_id_mat(norm)
    p_pos = np.empty((m, n), np.int64)
    np.put_along_axis(p_pos, p_ids, np.broadcast_to(np.arange(1, n + 1), (m, n)), 1)
    need = np.maximum(t_pos, p_pos)                 # item is in both top-k sets iff k >= need
    hits = np.zeros((m, n + 2), np.int64)
    rows = np.repeat(np.arange(m), n)
    np.add.at(hits, (rows, need.ravel()), 1)
    cum = np.cumsum(hits, axis=1)                   # cum[:, k] = |true_topk_incl(k) & pred_topk(k)|
    ka = np.asarray(ks)
    ps = cum[:, ka] / ka[None, :]
    for i in print_ids:
        print('query {}\nks:    {}\nprecs: {}'.format(i, ks, ps[i]))
    return np.mean(ps, axis=0)


def mean_reciprocal_rank(true_r, pred_r, norm, print_ids=[]):
    m, n = true_r.m_n()
    assert true_r.m_n() == pred_r.m_n()
    t_pos = _tie_inclusive_position(true_r.sort_id_mat(norm), true_r.dist_sim_mat(norm))
    p_rank = _tie_inclusive_position(pred_r.sort_id_mat(norm), pred_r.dist_sim_mat(norm))
    top = t_pos == 1                                 # tie set of the true best answer
    ranks = np.where(top, p_rank, n + 1).min(axis=1).astype(np.float64)
    for i in print_ids:
        print('query {}\nrank: {}'.format(i, ranks[i]))
    return float(np.mean(1.0 / ranks))              # == 1 / hmean(ranks)


def mean_squared_error(true_r, pred_r, sim_kernel, yeta, norm):
    m, n = true_r.m_n()
    assert true_r.m_n() == pred_r.m_n()
    return np.linalg.norm(true_r.sim_mat(sim_kernel, yeta, norm) -
                          pred_r.sim_mat(sim_kernel, yeta, norm)) / (m * n)


def average_time(r):
    return np.mean(r.time_mat())


def kendall_tau_counts(x, y):
    """Exact integer (concordant, discordant, ties only in x, ties only in y, joint ties) of
    one row pair in O(n log n) (Knight's algorithm: sort by x then y, count inversions in y
    with a merge sort, tie group sizes by run lengths)."""
    x = np.asarray(x)
    y = np.asarray(y)
    n = len(x)
    perm = np.lexsort((y, x))
    xs, ys = x[perm], y[perm]

    def _pairs_in_runs(new_run):
        idx = np.nonzero(np.append(new_run, True))[0]
        r = np.diff(idx).astype(np.int64)
        return int(np.sum(r * (r - 1) // 2))
    nr_x = np.ones(n, bool); nr_x[1:] = xs[1:] != xs[:-1]
    nr_xy = nr_x.copy(); nr_xy[1:] |= ys[1:] != ys[:-1]
    n_x, n_xy = _pairs_in_runs(nr_x), _pairs_in_runs(nr_xy)
    ysort = np.sort(y, kind='mergesort')
    nr_y = np.ones(n, bool); nr_y[1:] = ysort[1:] != ysort[:-1]
    n_y = _pairs_in_runs(nr_y)
    # discordant = strict inversions of ys (x ascending, ties in x sorted by y so they add none)
    rank = np.searchsorted(ysort, ys, side='left')
    dis = _count_inversions(rank)
    tot = n * (n - 1) // 2
    con = tot - n_x - n_y + n_xy - dis
    return con, dis, n_x - n_xy, n_y - n_xy, n_xy


def _count_inversions(a):
    """# of pairs i<j with a[i] > a[j] (bottom-up merge sort, numpy per level)."""
    a = np.asarray(a, np.int64).copy()
    n = len(a)
    inv = 0
    width = 1
    while width < n:
        for lo in range(0, n, 2 * width):
            mid, hi = min(lo + width, n), min(lo + 2 * width, n)
            if mid >= hi:
                continue
            left, right = a[lo:mid], a[mid:hi]
            # for each right element: # of left elements strictly greater
            inv += int(np.sum(len(left) - np.searchsorted(left, right, side='right')))
            a[lo:hi] = np.sort(a[lo:hi], kind='mergesort')
        width *= 2
    return inv


def kendall_tau(true_r, pred_r, norm):
    """Mean over queries of tau-b between the true and predicted SIMILARITY rows (a distance
    ground truth is negated so that higher = more similar on both sides)."""
    dev = _on_device(true_r, pred_r, norm)
    if dev is not None:
        taus = _ENGINE.kendall_tau(dev[0], dev[1])[0].cpu().numpy()
        # tau-b is undefined (NaN) for a row that is constant on either side; a sweep point whose model collapsed
        # to a constant has only such rows
        return (float(np.nanmean(taus)) if np.any(~np.isnan(taus)) else float('nan')), taus
    t = true_r.dist_sim_mat(norm)
    if hasattr(true_r, 'dist_mat'):
        t = -t
    p = pred_r.dist_sim_mat(norm)
    if hasattr(pred_r, 'dist_mat'):
        p = -p
    taus = []
    for i in range(t.shape[0]):
        con, dis, tx, ty, txy = kendall_tau_counts(t[i], p[i])
        den = np.sqrt(float(con + dis + tx)) * np.sqrt(float(con + dis + ty))
        taus.append((con - dis) / den if den > 0 else np.nan)
    return float(np.nanmean(taus)), np.array(taus)
