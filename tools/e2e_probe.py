"""Where the end-to-end ranking step's time goes: pinned H2D copy alone, kernels alone, rank_host."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from graphembedding_b200.engine import SgnnEngine
from graphembedding_b200.layers_factory import parse_layers
enc, q, db = bench.make_rank_data(0)[:3]
eng = SgnnEngine(parse_layers(bench.DEFAULT_LAYERS, 29), 29, precision='bf16x3')
eng.init_params(123)
eng.reserve(max_graphs=bench.N_DB, max_nodes=db.total_nodes, max_nnz=db.nnz, max_queries=bench.M_QUERIES, max_db=bench.N_DB)
dev = eng.device
qp, dbp = q.pin(), db.pin()
ids_h = torch.empty((bench.M_QUERIES, 10), dtype=torch.int32).pin_memory()
vals_h = torch.empty((bench.M_QUERIES, 10), dtype=torch.float32).pin_memory()
def t(fn, n=50):
    for _ in range(5): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e6
dst = torch.empty(dbp._arena.numel(), dtype=torch.uint8, device=dev)
print('pinned arena bytes', dbp._arena.numel())
print('H2D db arena copy (torch, sync each)   %.1f us' % t(lambda: (dst.copy_(dbp._arena, non_blocking=True), torch.cuda.synchronize())))
small = torch.empty(80000, dtype=torch.uint8).pin_memory(); sd = torch.empty(80000, dtype=torch.uint8, device=dev)
print('D2H 80 KB (sync each)                  %.1f us' % t(lambda: (small.copy_(sd, non_blocking=True), torch.cuda.synchronize())))
print('empty sync                             %.1f us' % t(lambda: torch.cuda.synchronize()))
qd, dbd = q.to(dev), db.to(dev)
emb_q = torch.empty((bench.M_QUERIES, eng.emb_dim), device=dev); emb_db = torch.empty((bench.N_DB, eng.emb_dim), device=dev)
sim = torch.empty((bench.M_QUERIES, bench.N_DB), device=dev)
def dev_step():
    eng.embed(qd, out=emb_q); eng.embed(dbd, out=emb_db); eng.score_all_pairs(emb_q, emb_db, out=sim); eng.topk(sim, 10)
print('device-resident step, async x50         %.1f us' % t(dev_step))
print('device-resident step, sync each         %.1f us' % t(lambda: (dev_step(), torch.cuda.synchronize())))
print('rank_host                               %.1f us' % t(lambda: eng.rank_host(qp, dbp, 10, ids=ids_h, vals=vals_h)))
print('rank_host (k=0: no topk, no D2H)        %.1f us' % t(lambda: eng.rank_host(qp, dbp, 0, ids=ids_h, vals=vals_h)))
