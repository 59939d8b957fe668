"""GPU parity of the tensor-core node-node histogram (csrc/sgnn_hist_tc.cuh, C ABI sgnn_hist_all_pairs) and of
the SimGNN-style all-pairs path built on it.

The Histogram layer is BUILD-DEFINED (absent from the reference snapshot, SURVEY a24 / F1: "parity unpinned by
reference"); the pin is oracle/siamese_oracle.py: pair_histogram.  Bin counts are integers and must be EQUAL to
the oracle's for every pair whose closest node-node score is farther than EDGE (in sigmoid units) from a bin
edge; cells inside that band may land in the neighbouring bin (the tensor-core dot product is a 6-product bf16
split accumulated in fp32, ~1e-7 relative) and such pairs are excluded explicitly, never more than a few %."""
import types

import numpy as np
import pytest

from util import CONFIGS, O, corpus, encoder_for, oracle_graphs

torch = pytest.importorskip('torch')
pytestmark = pytest.mark.gpu

EDGE = 1e-5
RTOL = 1e-4


def _engine(cfg_name, **kw):
    from graphembedding_b200.engine import SgnnEngine
    from graphembedding_b200.layers_factory import parse_layers
    layers = parse_layers(CONFIGS[cfg_name], 29)
    eng = SgnnEngine(layers, 29, **kw)
    flat = O.init_params(O.parse_layer_specs(CONFIGS[cfg_name], 29), seed=11)
    flat = flat + (np.random.default_rng(5).standard_normal(flat.shape) * 0.01).astype(np.float32)
    eng.set_params(flat)
    return eng, layers, flat


def _side(sizes, d, rng, scale, device):
    """A synthetic batch of node rows: sizes[g] rows of width d per graph."""
    off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    rows = (rng.standard_normal((int(off[-1]), d)) * scale).astype(np.float32)
    b = types.SimpleNamespace(n_graphs=len(sizes), node_off=torch.as_tensor(off, device=device),
                              max_nodes=int(max(sizes)), total_nodes=int(off[-1]))
    return b, rows, off


def _oracle_counts(rows_q, off_q, rows_db, off_db, bins):
    """(counts [m][n][bins], margin [m][n]) from the fp64 restatement of the layer (oracle pair_histogram)."""
    m, n = len(off_q) - 1, len(off_db) - 1
    with np.errstate(over='ignore'):          # exp overflows to inf for very negative scores: sigmoid = 0, as intended
        s = 1.0 / (1.0 + np.exp(-(rows_q.astype(np.float64) @ rows_db.astype(np.float64).T)))
    idx = np.minimum(bins - 1, np.floor(s * bins).astype(np.int64))
    mar = np.abs(s * bins - np.clip(np.round(s * bins), 1, bins - 1)) / bins     # distance to the nearest INTERIOR edge
    counts = np.zeros((m, n, bins), np.int64)
    margin = np.full((m, n), np.inf)
    for i in range(m):
        for j in range(n):
            blk = idx[off_q[i]:off_q[i + 1], off_db[j]:off_db[j + 1]]
            if blk.size:
                counts[i, j] = np.bincount(blk.reshape(-1), minlength=bins)
                margin[i, j] = mar[off_q[i]:off_q[i + 1], off_db[j]:off_db[j + 1]].min()
    return counts, margin


def test_oracle_counts_helper_is_the_oracles_pair_histogram():
    rng = np.random.default_rng(0)
    a, b = rng.standard_normal((5, 16)), rng.standard_normal((7, 16))
    _, want, _ = O.pair_histogram(torch.as_tensor(a), torch.as_tensor(b), 16)
    got, _ = _oracle_counts(a.astype(np.float32).astype(np.float64), [0, 5], b.astype(np.float32).astype(np.float64), [0, 7], 16)
    # (float32 rounding of the rows can move a cell only if it sits on an edge; the seeded draw has none)
    assert np.array_equal(got[0, 0], want)


@pytest.mark.parametrize('cfg,m,n,qmax,dbmax,scale', [
    ('simgnn', 37, 301, 10, 10, 0.5),      # AIDS-shaped
    ('simgnn', 64, 257, 15, 17, 1.0),      # the largest sizes one count byte holds (15 x 17 = 255 cells)
    ('simgnn', 9, 1500, 3, 25, 3.0),       # wide scores: the outer bins fill up, sigma == 1 / == 0 cells
    ('head_drop', 33, 280, 12, 12, 0.7),   # 7 bins, node width 12
])
def test_hist_all_pairs_counts_equal_the_oracle(cfg, m, n, qmax, dbmax, scale):
    eng, layers, flat = _engine(cfg, precision='bf16x3')
    d = eng.node_dim
    bins = [L['bins'] for L in layers if L['kind'] == 'Histogram'][0]
    rng = np.random.default_rng(m * 1000 + n)
    sq = rng.integers(1, qmax + 1, size=m); sq[0] = qmax; sq[-1] = 1
    sd = rng.integers(1, dbmax + 1, size=n); sd[0] = dbmax; sd[n // 2] = 1
    bq, rq, oq = _side(sq, d, rng, scale, eng.device)
    bd, rd, od = _side(sd, d, rng, scale, eng.device)
    rq[: sq[0]] = 0.0                      # an all-zero query graph: every cell sits exactly ON the sigmoid(0) edge
    if scale >= 3.0:
        rd[:4] *= 40.0                     # |s| far beyond the clamp: bins 0 and B-1
    eng.reserve(max_nodes=int(od[-1]), max_queries=m, max_db=n)
    got = eng.hist_all_pairs(torch.as_tensor(rq, device=eng.device), bq, torch.as_tensor(rd, device=eng.device), bd)
    torch.cuda.synchronize()
    got = got.cpu().numpy().astype(np.int64)
    want, margin = _oracle_counts(rq, oq, rd, od, bins)
    assert np.all(got[:, :, bins:] == 0)
    # every pair accounts for exactly its cells, in bins that exist
    assert np.array_equal(got[:, :, :bins].sum(-1), np.outer(sq, sd))
    ok = margin > EDGE
    ok[0, :] = True                        # the zero query: sigmoid(0) = 1/2 is an exact edge and must go UP (bin floor(B/2))
    assert ok.mean() > 0.97
    assert np.array_equal(got[:, :, :bins][ok], want[ok])
    assert np.all(got[0, :, bins // 2] == sq[0] * sd)
    # pairs inside the band differ by at most the cells that sit in it
    diff = np.abs(got[:, :, :bins] - want).sum(-1)
    assert diff.max() <= 4


def test_hist_all_pairs_rejects_sizes_outside_the_kernel():
    eng, layers, flat = _engine('simgnn', precision='bf16x3')
    rng = np.random.default_rng(1)
    bq, rq, oq = _side([16, 3], 16, rng, 1.0, eng.device)       # 16 > 15 query nodes
    bd, rd, od = _side([5, 5, 5], 16, rng, 1.0, eng.device)
    eng.reserve(max_nodes=64, max_queries=2, max_db=3)
    with pytest.raises(ValueError):
        eng.hist_all_pairs(torch.as_tensor(rq, device=eng.device), bq, torch.as_tensor(rd, device=eng.device), bd)
    eng32, _, _ = _engine('simgnn', precision='fp32')
    with pytest.raises(RuntimeError):      # the CUDA-core precision has no tensor-core histogram
        eng32.hist_all_pairs(torch.as_tensor(rq, device=eng.device), bq, torch.as_tensor(rd, device=eng.device), bd)


@pytest.mark.parametrize('cfg,n_hi', [('simgnn', 10), ('head_drop', 10), ('simgnn', 15)])
def test_all_pairs_with_tensor_core_histogram_matches_oracle_and_cuda_core_path(cfg, n_hi):
    """The m x n loop of test() (train.py:57-61) on the SimGNN-style chain: tensor-core histogram + head kernel
    against the fp64 oracle on EVERY cell outside the bin-edge band, and against the CUDA-core (fp32) engine."""
    from graphembedding_b200.graphs import pack_nx_graphs
    eng, layers, flat = _engine(cfg, precision='bf16x3', final_act='sigmoid')
    eng32, _, _ = _engine(cfg, precision='fp32', final_act='sigmoid')
    gs = corpus(190, 2, n_hi, 3)
    enc = encoder_for(gs)
    m, n = 29, 161
    qb, db = pack_nx_graphs(gs[:m], enc), pack_nx_graphs(gs[m:], enc)
    sims = []
    for e in (eng, eng32):
        e.reserve(max_nodes=qb.total_nodes + db.total_nodes, max_queries=m, max_db=n)
        qd, dd = qb.to(e.device), db.to(e.device)
        eq, rq = e.embed_nodes(qd)
        ed, rd = e.embed_nodes(dd)
        sims.append(e.score_all_pairs_nodes(eq, rq, qd, ed, rd, dd).cpu().numpy())
    P = O.unflatten(layers, flat, torch.float64)
    og = oracle_graphs(gs, enc)
    g1s = [og[i] for i in range(m) for _ in range(n)]
    g2s = [og[m + j] for _ in range(m) for j in range(n)]
    ok = (O.pair_bin_margins(layers, P, g1s, g2s, 29) > EDGE).reshape(m, n)
    assert ok.mean() > 0.99
    ref = torch.sigmoid(O.forward_pairs(layers, P, g1s, g2s, 29, torch.float64)).numpy().reshape(m, n)
    assert np.max(np.abs(sims[0] - ref)[ok]) <= RTOL
    assert np.max(np.abs(sims[1] - ref)[ok]) <= RTOL
    assert np.max(np.abs(sims[0] - sims[1])[ok]) <= RTOL


def test_tensor_core_engine_falls_back_when_the_database_graphs_are_too_large():
    """Queries fit the tensor-core histogram (<= 10 nodes) but the database does not (30-node graphs: 10 x 30 cells do
    not fit a count byte): the scorer stage takes the CUDA-core kernels, building the per-query tables it skipped on the
    query branch.  Same scores as the fp32 engine, through the staged ranking entry point too."""
    from graphembedding_b200.graphs import pack_nx_graphs
    eng, layers, flat = _engine('simgnn', precision='bf16x3', final_act='sigmoid')
    eng32, _, _ = _engine('simgnn', precision='fp32', final_act='sigmoid')
    qs, dbs = corpus(20, 2, 10, 5), corpus(90, 12, 30, 6)
    enc = encoder_for(qs + dbs)
    qb, db = pack_nx_graphs(qs, enc), pack_nx_graphs(dbs, enc)
    assert db.max_nodes > 25
    out = []
    for e in (eng, eng32):
        e.reserve(max_graphs=90, max_nodes=qb.total_nodes + db.total_nodes, max_nnz=db.nnz, max_queries=20, max_db=90)
        qd, dd = qb.to(e.device), db.to(e.device)
        eq, rq = e.embed_nodes(qd)
        ed, rd = e.embed_nodes(dd)
        a = e.score_all_pairs_nodes(eq, rq, qd, ed, rd, dd).cpu().numpy()
        ids, vals, sim = e.rank_device(qd, dd, 5)
        assert np.array_equal(sim.cpu().numpy(), a)
        out.append(a)
    # both engines scored with the same CUDA-core kernels; the embeddings differ at rounding level (the tensor-core engine
    # runs its dense GCN layers as 6-product bf16 GEMMs)
    assert np.max(np.abs(out[0] - out[1])) <= 1e-6
