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
enc, q, db = bench.make_rank_data(rank)[:3]
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

# ---- data-parallel training: peer-memory gradient push fused into Adam vs NCCL all-reduce + Adam ----
def train_run(use_peer, steps=6):
    enc2, batch, batches = bench.make_train_data(rank, world)
    e = SgnnEngine(parse_layers(bench.TRAIN_LAYERS, 29), 29, device=local, loss_mode='paired', ntn_mode='reference')
    e.init_params(123)
    pg = DS.PeerTopk(e, 0, 0, dist) if use_peer else None
    bd = batch.to(e.device)
    feeds = [(torch.as_tensor(a, device=e.device), torch.as_tensor(b, device=e.device),
              torch.as_tensor(t, device=e.device), s1, s2) for a, b, t, s1, s2 in batches]
    loss = torch.zeros(4, dtype=torch.float32, device=e.device)
    gb = bench.TRAIN_BATCH * world
    def step(i):
        g1, g2, t, s1, s2 = feeds[i % len(feeds)]
        loss.zero_()
        DS.data_parallel_step(e, bd, g1, g2, t, gb, s1, s2, loss, dist, peer=pg)
    for i in range(steps):
        step(i)
    torch.cuda.synchronize()
    params = e.get_params()
    dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(200):
        step(i)
    b.record()
    torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b) / 200], device=e.device); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if pg is not None:
        pg.close()
    return params, float(t.item())

p_peer, ms_peer = train_run(True)
p_nccl, ms_nccl = train_run(False)
# identical on every rank (rank-ordered sum), and equal to the NCCL path up to the rounding of unordered atomics
gathered = [None] * world
dist.all_gather_object(gathered, p_peer)
assert all(np.array_equal(gathered[0], g) for g in gathered), 'ranks diverged'
err = float(np.max(np.abs(p_peer - p_nccl)))
assert err <= 1e-6, err   # gradient atomics are unordered: runs differ at rounding level
if rank == 0:
    print('dp train OK: params max |p2p - nccl| = %.3g; step ms  p2p %.4f  nccl %.4f' % (err, ms_peer, ms_nccl))
dist.barrier()
dist.destroy_process_group()
