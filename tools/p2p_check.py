"""2+ GPU check of the peer-memory top-k exchange (run under torchrun):
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/p2p_check.py
Each rank ranks its database shard; the P2P exchange must equal the all-gather + merge path and the
top-k of the concatenated full similarity matrix; then both paths are timed."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
import bench
from graphembedding_b200 import distributed as DS
from graphembedding_b200.engine import SgnnEngine
from graphembedding_b200.layers_factory import parse_layers

rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local)
dist.init_process_group('nccl', device_id=torch.device('cuda', local))
torch.cuda.set_stream(torch.cuda.Stream())
enc, q, db = bench.make_rank_data(rank)
eng = SgnnEngine(parse_layers(bench.DEFAULT_LAYERS, 29), 29, device=local, precision='bf16x3')
eng.init_params(123)
eng.reserve(max_graphs=bench.N_DB, max_nodes=db.total_nodes, max_nnz=db.nnz, max_queries=bench.M_QUERIES, max_db=bench.N_DB)
qd, dbd = q.to(eng.device), db.to(eng.device)
K = 10
peer = DS.PeerTopk(eng, bench.M_QUERIES, K, dist)
for it in range(4):
    ids_p, vals_p = DS.sharded_rank(eng, qd, dbd, rank * bench.N_DB, K, dist, peer=peer)
    ids_n, vals_n = DS.sharded_rank(eng, qd, dbd, rank * bench.N_DB, K, dist)
    torch.cuda.synchronize()
    assert torch.equal(ids_p, ids_n) and torch.equal(vals_p, vals_n), (rank, it)
# against the full matrix: gather every rank's sim tile
_, _, sim = eng.rank_device(qd, dbd, 0)
tiles = [torch.empty_like(sim) for _ in range(world)]
dist.all_gather(tiles, sim.contiguous())
full = torch.cat(tiles, 1)
ids_f, vals_f = eng.topk(full, K)
torch.cuda.synchronize()
assert torch.equal(ids_f, ids_p) and torch.equal(vals_f, vals_p), rank
flush = torch.empty(256 << 20, dtype=torch.uint8, device=eng.device)
res = {}
for name, pr in (('p2p', peer), ('nccl', None)):
    for _ in range(5):
        DS.sharded_rank(eng, qd, dbd, rank * bench.N_DB, K, dist, peer=pr)
    torch.cuda.synchronize(); dist.barrier()
    ms = bench.timed_steps(torch, lambda: DS.sharded_rank(eng, qd, dbd, rank * bench.N_DB, K, dist, peer=pr), 30, flush) / 30
    t = torch.tensor([ms], device=eng.device); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    res[name] = float(t.item())
if rank == 0:
    print('p2p_check OK: world %d; step ms  p2p %.4f  nccl %.4f' % (world, res['p2p'], res['nccl']))
peer.close()
dist.barrier()
dist.destroy_process_group()
