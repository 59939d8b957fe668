"""Random-shape sweep of the tensor-core histogram kernels against the fp64 statement (tests/test_gpu_hist_tc.py:
_oracle_counts; tests/test_gpu_hist_pairs_tc.py: _oracle): sizes at and around every tile / batch / bucket boundary."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tests'))
import numpy as np, torch
import test_gpu_hist_tc as T
import test_gpu_hist_pairs_tc as TP

rng = np.random.default_rng(int(os.environ.get('FUZZ_SEED', '1')))
bad = 0
for case in range(int(os.environ.get('FUZZ_CASES', '40'))):
    cfg = 'simgnn' if case % 3 else 'head_drop'
    eng, layers, flat = T._engine(cfg, precision='bf16x3')
    d = eng.node_dim
    bins = [L['bins'] for L in layers if L['kind'] == 'Histogram'][0]
    qmax = int(rng.integers(1, 16)); dbmax = int(rng.integers(1, min(64, 255 // qmax) + 1))
    m = int(rng.choice([1, 2, 7, 8, 9, 63, 64, 65, 127, 128, 129, 200, 300])); n = int(rng.choice([1, 5, 127, 128, 129, 255, 256, 257, 1000, 2049]))
    sq = rng.integers(1, qmax + 1, size=m); sd = rng.integers(1, dbmax + 1, size=n)
    sq[rng.integers(0, m)] = qmax; sd[rng.integers(0, n)] = dbmax
    scale = float(rng.choice([0.3, 1.0, 3.0]))
    bq, rq, oq = T._side(sq, d, rng, scale, eng.device)
    bd, rd, od = T._side(sd, d, rng, scale, eng.device)
    eng.reserve(max_nodes=int(max(od[-1], oq[-1])), max_queries=m, max_db=n)
    got = eng.hist_all_pairs(torch.as_tensor(rq, device=eng.device), bq, torch.as_tensor(rd, device=eng.device), bd)
    torch.cuda.synchronize()
    got = got.cpu().numpy().astype(np.int64)
    want, margin = T._oracle_counts(rq, oq, rd, od, bins)
    ok = margin > T.EDGE
    sums = np.array_equal(got[:, :, :bins].sum(-1), np.outer(sq, sd)) and np.all(got[:, :, bins:] == 0)
    eq = np.array_equal(got[:, :, :bins][ok], want[ok])
    dmax = int(np.abs(got[:, :, :bins] - want).sum(-1).max())
    good = sums and eq and dmax <= 4 and (ok.mean() > 0.9 or m * n < 100)
    bad += not good
    print('all-pairs', case, cfg, 'm', m, 'n', n, 'qmax', qmax, 'dbmax', dbmax, 'scale', scale, 'ok' if good else 'FAIL', dmax, round(float(ok.mean()), 3), flush=True)
for case in range(int(os.environ.get('FUZZ_PAIR_CASES', '12'))):
    cfg = 'simgnn' if case % 2 else 'head_drop'
    eng = TP._engine(cfg, 'bf16x3')
    d = eng.node_dim; bins = 16 if cfg == 'simgnn' else 7
    G = int(rng.integers(2, 20))
    sizes = rng.choice([1, 31, 32, 33, 100, 127, 128, 129, 255, 256, 257, 400, 511, 512, 513], size=G)
    if sizes.sum() < 48 * G:
        sizes[0] = 48 * G
    off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    rows = (rng.standard_normal((int(off[-1]), d)) * float(rng.choice([0.3, 1.0]))).astype(np.float32)
    P = int(rng.integers(1, 60))
    g1 = rng.integers(0, G, P).astype(np.int32); g2 = rng.integers(0, G, P).astype(np.int32)
    dev = eng.device
    hist, counts = eng.pair_histogram(torch.as_tensor(rows, device=dev), torch.as_tensor(off, device=dev),
                                      torch.as_tensor(g1, device=dev), torch.as_tensor(g2, device=dev), bins=bins)
    torch.cuda.synchronize()
    got = counts.cpu().numpy().astype(np.int64)
    want, band = TP._oracle(rows, off, g1, g2, bins)
    diff = np.abs(got - want).sum(1)
    good = np.array_equal(got.sum(1), sizes[g1] * sizes[g2]) and bool(np.all(diff <= 2 * band))
    bad += not good
    print('pairs', case, cfg, 'G', G, 'P', P, 'max nodes', int(sizes.max()), 'ok' if good else 'FAIL', int(diff.max()), int(band.max()), flush=True)
print('FUZZ', 'OK' if bad == 0 else 'FAILED %d' % bad)
