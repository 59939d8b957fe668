"""Times the stages of the SimGNN-style ranking step with the tensor-core histogram on the bench data
(1 000 queries x 10 000 database graphs) and compares its scores with the CUDA-core (fp32) engine.
  python tools/hist_probe.py            # timings + cross-check
  SGNN_NO_GRAPHS=1 PROBE_NCU=1 ncu ... python tools/hist_probe.py   # a few plain launches for ncu"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import bench
from graphembedding_b200.engine import SgnnEngine
from graphembedding_b200.layers_factory import parse_layers

M, N = int(os.environ.get('PROBE_M', bench.M_QUERIES)), int(os.environ.get('PROBE_N', bench.N_DB))
bench.M_QUERIES, bench.N_DB = M, N
enc, q, db = bench.make_rank_data(0)[:3]


def engine(precision):
    e = SgnnEngine(parse_layers(bench.SIMGNN_LAYERS, 29), 29, final_act='sigmoid', loss_mode='paired', precision=precision)
    e.init_params(123)
    e.reserve(max_graphs=N, max_nodes=max(q.total_nodes, db.total_nodes), max_nnz=db.nnz, max_queries=M, max_db=N)
    return e


eng = engine('bf16x3')
qd, dbd = q.to(eng.device), db.to(eng.device)
eq, rq = eng.embed_nodes(qd)
ed, rd = eng.embed_nodes(dbd)
sim = torch.empty((M, N), device=eng.device)
cnt = torch.zeros((M, N, 16), dtype=torch.uint8, device=eng.device)
torch.cuda.synchronize()


def timed(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


if os.environ.get('PROBE_NCU'):
    for _ in range(2):
        eng.hist_all_pairs(rq, qd, rd, dbd, out=cnt)
        eng.score_all_pairs_nodes(eq, rq, qd, ed, rd, dbd, out=sim)
    torch.cuda.synchronize()
    print('ok')
    sys.exit(0)

out = {'m': M, 'n': N, 'cells': float(q.total_nodes) * float(db.total_nodes)}
out['hist_prep_q_ms'] = timed(lambda: eng.hist_all_pairs(rq, qd, rd, dbd, out=cnt, stages=1))
out['hist_prep_db_ms'] = timed(lambda: eng.hist_all_pairs(rq, qd, rd, dbd, out=cnt, stages=2))
out['hist_main_ms'] = timed(lambda: eng.hist_all_pairs(rq, qd, rd, dbd, out=cnt, stages=4))
out['score_all_pairs_nodes_ms'] = timed(lambda: eng.score_all_pairs_nodes(eq, rq, qd, ed, rd, dbd, out=sim))
out['rank_device_ms'] = timed(lambda: eng.rank_device(qd, dbd, 10))
c = cnt.cpu().numpy()
cells_pair = np.outer(np.diff(q.node_off), np.diff(db.node_off))
out['count_sums_ok'] = bool(np.array_equal(c.sum(-1), cells_pair))
# cross-check against the CUDA-core engine on a slice of the queries
eng32 = engine('fp32')
mq = min(M, 64)
from graphembedding_b200.graphs import pack_nx_graphs  # noqa
e32q, r32q = eng32.embed_nodes(qd)
e32d, r32d = eng32.embed_nodes(dbd)
sim32 = eng32.score_all_pairs_nodes(e32q, r32q, qd, e32d, r32d, dbd)
out['fp32_head_kernel_ms'] = timed(lambda: eng32.score_all_pairs_nodes(e32q, r32q, qd, e32d, r32d, dbd, out=sim32), reps=3)
d = (sim - sim32).abs()
out['max_abs_diff_vs_cuda_core'] = float(d.max())
out['frac_cells_diff_gt_1e-4'] = float((d > 1e-4).float().mean())
print(json.dumps(out))
