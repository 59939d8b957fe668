"""Runs the SimGNN-style ranking step (embed with node rows, all-pairs head kernel, top-k) and one training
step a few times, for ncu:
  SGNN_NO_GRAPHS=1 ncu --set full -k regex:allpairs_head -c 1 -o gpurun_out/head python tools/head_probe.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from graphembedding_b200.engine import SgnnEngine
from graphembedding_b200.layers_factory import parse_layers
enc, q, db = bench.make_rank_data(0)[:3]
eng = SgnnEngine(parse_layers(bench.SIMGNN_LAYERS, 29), 29, final_act='sigmoid', loss_mode='paired')
eng.init_params(123)
eng.reserve(max_graphs=bench.N_DB, max_nodes=db.total_nodes, max_nnz=db.nnz, max_queries=bench.M_QUERIES, max_db=bench.N_DB)
qd, dbd = q.to(eng.device), db.to(eng.device)
for _ in range(int(os.environ.get('PROBE_STEPS', '2'))):
    ids, vals, sim = eng.rank_device(qd, dbd, 10)
torch.cuda.synchronize()
_, batch, batches = bench.make_train_data(0, 1, n_batches=1)
bd = batch.to(eng.device)
a, b, t, s1, s2 = batches[0]
g1, g2, tt = (torch.as_tensor(v, device=eng.device) for v in (a, b, t))
for _ in range(2):
    loss = eng.train_step(bd, g1, g2, tt)
    eng.apply_adam(1.0, loss)
torch.cuda.synchronize()
print('ok', float(sim.mean()), float(loss[0]))
