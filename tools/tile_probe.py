"""Where the small-graph embedding kernel's time goes: launch-size sweep (fixed vs per-graph cost,
warm and flushed L2) and, with the instrumented library (SGNN_TILE_TRACE), per-phase clock stamps
of warp 0's first tile of every CTA.
  SGNN_DEFINES=-DSGNN_TILE_TRACE SGNN_LIB_SUFFIX=_trace python graphembedding_b200/csrc/build.py   # here
  SGNN_LIB_SUFFIX=_trace python tools/tile_probe.py                                                 # on the B200
"""
import ctypes as C
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import bench
from graphembedding_b200 import native
from graphembedding_b200.engine import SgnnEngine
from graphembedding_b200.layers_factory import parse_layers

enc, q, db = bench.make_rank_data(0)[:3]
eng = SgnnEngine(parse_layers(bench.DEFAULT_LAYERS, 29), 29, precision='bf16x3')
eng.init_params(123)
eng.reserve(max_graphs=bench.N_DB, max_nodes=db.total_nodes, max_nnz=db.nnz, max_queries=bench.M_QUERIES,
            max_db=bench.N_DB)
dev = eng.device
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timed(fn, cold, n=20):
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
    for _ in range(3):
        fn()
    for a, b in ev:
        if cold:
            flush.fill_(1)
        a.record(); fn(); b.record()
    torch.cuda.synchronize()
    return float(np.median([a.elapsed_time(b) for a, b in ev])) * 1e3


for name, batch in (('q (1000 graphs)', q), ('db (10000 graphs)', db)):
    bd = batch.to(dev)
    out = torch.empty((batch.n_graphs, eng.emb_dim), device=dev)
    for cold in (False, True):
        print('%-18s %s L2: %.1f us' % (name, 'cold' if cold else 'warm', timed(lambda: eng.embed(bd, out=out), cold)))
    print('  total nodes %d  nnz %d  max_nodes %d' % (batch.total_nodes, batch.nnz, batch.max_nodes))

if os.environ.get('SGNN_LIB_SUFFIX'):
    fn = native.lib().sgnn_debug_tile_trace
    fn.restype = C.c_int
    fn.argtypes = [C.c_void_p, C.c_int]
    for name, batch in (('q', q), ('db', db)):
        bd = batch.to(dev)
        out = torch.empty((batch.n_graphs, eng.emb_dim), device=dev)
        for cold in (False, True):
            if cold:
                flush.fill_(1)
            torch.cuda.synchronize()
            eng.embed(bd, out=out)
            tr = np.zeros(1024 * 8, np.uint64)
            assert fn(tr.ctypes.data, tr.size) == 0
            tr = tr.reshape(1024, 8).astype(np.int64)
            tr = tr[tr[:, 0] > 0]
            t0 = tr[:, 0].min()
            print('%s %s: %d CTAs; start spread %.2f us; kernel span %.2f us' % (
                name, 'cold' if cold else 'warm', len(tr), (tr[:, 0].max() - t0) / 1e3, (tr[:, 7].max() - t0) / 1e3))
            ph = ['stage weights', 'tile setup (CSR)', 'node layers', 'pooling', 'remaining tiles']
            for i, nm in enumerate(ph):
                d = tr[:, i + 2] - tr[:, i + 1]
                print('   %-18s median %7d cyc  max %7d cyc' % (nm, int(np.median(d)), int(d.max())))
            print('   CTA lifetime (globaltimer) median %.2f us max %.2f us' % (
                float(np.median(tr[:, 7] - tr[:, 0])) / 1e3, float((tr[:, 7] - tr[:, 0]).max()) / 1e3))
            eng_dummy = None
