"""Runs the device-resident ranking step a few times (for an ncu launch list):
  ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/x.csv python tools/rank_probe.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from graphembedding_b200.engine import SgnnEngine
from graphembedding_b200.layers_factory import parse_layers
enc, q, db = bench.make_rank_data(0)[:3]
eng = SgnnEngine(parse_layers(bench.DEFAULT_LAYERS, 29), 29, precision='bf16x3')
eng.init_params(123)
eng.reserve(max_graphs=bench.N_DB, max_nodes=db.total_nodes, max_nnz=db.nnz, max_queries=bench.M_QUERIES, max_db=bench.N_DB)
qd, dbd = q.to(eng.device), db.to(eng.device)
for _ in range(int(os.environ.get('PROBE_STEPS', '3'))):
    ids, vals, sim = eng.rank_device(qd, dbd, 10)
torch.cuda.synchronize()
ids2, vals2 = eng.topk(sim, 10)
torch.cuda.synchronize()
print('match', bool(torch.equal(ids, ids2)))
# ---- check the slice hints against keys recomputed from sim ----
import ctypes as C
import numpy as np
from graphembedding_b200 import native
ptr, cap = C.c_void_p(), C.c_int64()
native.lib().sgnn_debug_slice_hints(eng.h, C.byref(ptr), C.byref(cap))
m, n = sim.shape
ns = ((n + 127) // 128) * 4
buf = (C.c_uint64 * (m * ns))()
torch.cuda.synchronize()
import ctypes
cudart = ctypes.CDLL('libcudart.so')
cudart.cudaMemcpy(buf, ptr, C.c_size_t(m * ns * 8), 2)
hints = np.frombuffer(buf, np.uint64).reshape(m, ns)
s = sim.cpu().numpy()
pad = np.full((m, ns * 32), -np.inf, np.float32); pad[:, :n] = s
u = (pad + 0.0).view(np.uint32)
u = np.where(u & 0x80000000, ~u, u | 0x80000000).astype(np.uint64)
keys = (u << np.uint64(32)) | np.arange(ns * 32, dtype=np.uint64)[None, :]
exp = keys.reshape(m, ns, 32).max(2)
bad = np.argwhere(hints != exp)
print('hint mismatches', len(bad), 'of', hints.size)
for q_, s_ in bad[:5]:
    print(q_, s_, hex(int(hints[q_, s_])), hex(int(exp[q_, s_])))
srt = np.sort(keys[:, :n], axis=1)[:, ::-1]
print('distinct top values in row 0:', len(np.unique(s[0])), 'max count', (s[0] == s[0].max()).sum())
if os.environ.get('SGNN_LIB_SUFFIX'):
    fn = native.lib().sgnn_debug_topk_trace
    fn.restype = C.c_int; fn.argtypes = [C.c_void_p, C.c_int]
    ids, vals, sim = eng.rank_device(qd, dbd, 10)
    dbg = np.zeros(1000 * 4, np.int32)
    assert fn(dbg.ctypes.data, dbg.size) == 0
    dbg = dbg.reshape(-1, 4)
    print('candidates: median %d max %d | hot slices: median %d max %d | cycles to candidates: median %d | fell back: %d rows' % (
        np.median(dbg[:, 0]), dbg[:, 0].max(), np.median(dbg[:, 1]), dbg[:, 1].max(), np.median(dbg[:, 2]), dbg[:, 3].sum()))
