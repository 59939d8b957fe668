"""Debug: mismatch statistics of sgnn_hist_all_pairs against the fp64 restatement on synthetic node rows."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tests'))
import numpy as np, torch
import test_gpu_hist_tc as T
for cfg, m, n, qmax, dbmax, scale in [('simgnn', 37, 301, 10, 10, 0.5), ('simgnn', 64, 257, 15, 17, 1.0),
                                      ('simgnn', 9, 1500, 3, 25, 3.0), ('head_drop', 33, 280, 12, 12, 0.7)]:
    eng, layers, flat = T._engine(cfg, precision='bf16x3')
    d = eng.node_dim
    bins = [L['bins'] for L in layers if L['kind'] == 'Histogram'][0]
    rng = np.random.default_rng(m * 1000 + n)
    sq = rng.integers(1, qmax + 1, size=m); sq[0] = qmax; sq[-1] = 1
    sd = rng.integers(1, dbmax + 1, size=n); sd[0] = dbmax; sd[n // 2] = 1
    bq, rq, oq = T._side(sq, d, rng, scale, eng.device)
    bd, rd, od = T._side(sd, d, rng, scale, eng.device)
    rq[: sq[0]] = 0.0
    if scale >= 3.0:
        rd[:4] *= 40.0
    eng.reserve(max_nodes=int(od[-1]), max_queries=m, max_db=n)
    got = eng.hist_all_pairs(torch.as_tensor(rq, device=eng.device), bq, torch.as_tensor(rd, device=eng.device), bd)
    torch.cuda.synchronize()
    got = got.cpu().numpy().astype(np.int64)
    want, margin = T._oracle_counts(rq, oq, rd, od, bins)
    diff = np.abs(got[:, :, :bins] - want).sum(-1)
    bad = np.argwhere(diff > 0)
    print(cfg, m, n, 'mismatching pairs', len(bad), 'of', m * n, 'sum ok', np.array_equal(got[:, :, :bins].sum(-1), np.outer(sq, sd)),
          'beyond bins', int(got[:, :, bins:].sum()))
    mb = margin[diff > 0]
    if len(bad):
        print('  margins of mismatching pairs: min %.3g median %.3g max %.3g; diff max %d' % (mb.min(), np.median(mb), mb.max(), diff.max()))
        print('  mismatching with margin > 1e-5:', int((mb > 1e-5).sum()), ' rows(q) hist', np.bincount(bad[:, 0], minlength=m)[:12])
        for i, j in bad[:5]:
            print('  pair', i, j, 'sizes', sq[i], sd[j], 'margin %.3g' % margin[i, j], 'got', got[i, j, :bins], 'want', want[i, j])
