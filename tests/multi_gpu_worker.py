"""Worker of tests/test_gpu_multi.py (run under torch.distributed.run, one process per GPU): the multi-GPU data
plane of SURVEY 8(e) on real peer memory.
  ranking : sgnn_rank_sharded (local ranking + peer push + merge, one CUDA graph) == all-gather + sgnn_topk_merge
            == top-k of the concatenated similarity matrix, BIT FOR BIT, over EPOCHS consecutive steps (the exchange
            slots are parity double-buffered: consecutive epochs must not alias);
  training: sgnn_apply_adam_allreduce leaves bit-identical parameters on every rank and stays within 1e-6 of the
            NCCL all-reduce + sgnn_apply_adam path."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
import numpy as np
import torch
import torch.distributed as dist

from util import CONFIGS, corpus, encoder_for
from graphembedding_b200 import distributed as DS
from graphembedding_b200.engine import SgnnEngine
from graphembedding_b200.graphs import pack_nx_graphs
from graphembedding_b200.layers_factory import parse_layers

EPOCHS = int(os.environ.get('MULTI_EPOCHS', '60'))
rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local)
dev = torch.device('cuda', local)
dist.init_process_group('nccl', device_id=dev)
torch.cuda.set_stream(torch.cuda.Stream())
M, N, K = 150, 1300, 10
gs = corpus(M + N * world, 2, 10, seed=77)
enc = encoder_for(gs)
q = pack_nx_graphs(gs[:M], enc)
db = pack_nx_graphs(gs[M + rank * N:M + (rank + 1) * N], enc)

for cfg, kw in (('default', {}), ('simgnn', dict(final_act='sigmoid'))):
    eng = SgnnEngine(parse_layers(CONFIGS[cfg], 29), 29, device=local, precision='bf16x3', **kw)
    eng.init_params(5)
    eng.reserve(max_graphs=N, max_nodes=max(q.total_nodes, db.total_nodes), max_nnz=db.nnz, max_queries=M, max_db=N)
    qd, dbd = q.to(dev), db.to(dev)
    peer = DS.PeerTopk(eng, M, K, dist)
    # the reference answer: top-k of the concatenated full matrix
    _, _, sim = eng.rank_device(qd, dbd, 0)
    tiles = [torch.empty_like(sim) for _ in range(world)]
    dist.all_gather(tiles, sim.contiguous())
    ids_f, vals_f = eng.topk(torch.cat(tiles, 1).contiguous(), K)
    ids_n, vals_n = DS.sharded_rank(eng, qd, dbd, rank * N, K, dist)          # all-gather + merge kernel
    torch.cuda.synchronize()
    assert torch.equal(ids_n, ids_f) and torch.equal(vals_n, vals_f), (cfg, rank, 'all-gather path')
    bufs = dict(emb_q=torch.empty((M, eng.emb_dim), device=dev), emb_db=torch.empty((N, eng.emb_dim), device=dev),
                sim=torch.empty((M, N), device=dev), ids=torch.empty((M, K), dtype=torch.int32, device=dev),
                vals=torch.empty((M, K), device=dev), out_ids=torch.empty((M, K), dtype=torch.int32, device=dev),
                out_vals=torch.empty((M, K), device=dev))
    for ep in range(EPOCHS):    # same buffers every step: eager, capture, then graph replays; no host sync between ranks
        bufs['out_ids'].zero_()
        ids_p, vals_p = DS.sharded_rank(eng, qd, dbd, rank * N, K, dist, peer=peer, **bufs)
        if ep % 7 == 0 or ep == EPOCHS - 1:
            torch.cuda.synchronize()
            assert torch.equal(ids_p, ids_f) and torch.equal(vals_p, vals_f), (cfg, rank, ep)
    torch.cuda.synchronize()
    assert torch.equal(ids_p, ids_f) and torch.equal(vals_p, vals_f), (cfg, rank, 'last')
    peer.close()
    # ranking metrics that need WHOLE rows (Kendall tau, precision@k) under database sharding: all-to-all of the local
    # [M, N] tiles into M / world full rows per rank, the metric kernels on those, results combined
    full = torch.cat(tiles, 1).contiguous()
    g = torch.Generator(device='cpu').manual_seed(5)
    true_full = (full.cpu() + 0.05 * torch.randn(full.shape, generator=g)).to(dev).contiguous()   # same on every rank
    prec, tau = DS.sharded_full_row_metrics(eng, true_full[:, rank * N:(rank + 1) * N].contiguous(), sim.contiguous(),
                                            [1, 10, 100], dist)
    ref_p = eng.precision_at_ks(true_full, full, [1, 10, 100]).cpu().numpy()
    ref_t = eng.kendall_tau(true_full, full)[0].cpu().numpy()
    assert np.array_equal(prec, ref_p) or np.allclose(prec, ref_p, rtol=0, atol=1e-12), (cfg, rank, prec, ref_p)
    assert np.array_equal(tau, ref_t), (cfg, rank, 'tau')
    if rank == 0:
        print('ranking ok:', cfg, 'world', world, 'epochs', EPOCHS, '+ sharded full-row metrics', flush=True)


def train_run(cfg, use_peer, steps=12):
    tg = corpus(256, 2, 20, seed=9)
    tb = pack_nx_graphs(tg, encoder_for(tg))
    kw = dict(final_act='sigmoid') if cfg == 'simgnn' else {}
    e = SgnnEngine(parse_layers(CONFIGS[cfg], 29), 29, device=local, loss_mode='paired', **kw)
    e.init_params(11)
    pg = DS.PeerTopk(e, 0, 0, dist) if use_peer else None
    bd = tb.to(dev)
    rng = np.random.default_rng(3)
    B = 32 * world
    loss = torch.zeros(4, dtype=torch.float32, device=dev)
    for i in range(steps):
        pairs = rng.integers(0, 256, size=(B, 2)).astype(np.int32)
        t = rng.uniform(0.1, 1.0, B).astype(np.float32)
        mine = slice(rank, None, world)
        s1, s2 = DS.global_target_stats(t)
        loss.zero_()
        DS.data_parallel_step(e, bd, torch.as_tensor(pairs[mine, 0], device=dev), torch.as_tensor(pairs[mine, 1], device=dev),
                              torch.as_tensor(t[mine], device=dev), B, s1, s2, loss, dist, peer=pg)
    torch.cuda.synchronize()
    p = e.get_params()
    if pg is not None:
        pg.close()
    return p


for cfg in ('avg_ntn', 'simgnn'):
    p_peer, p_nccl = train_run(cfg, True), train_run(cfg, False)
    got = [None] * world
    dist.all_gather_object(got, p_peer)
    assert all(np.array_equal(got[0], g) for g in got), (cfg, 'ranks diverged')
    err = float(np.max(np.abs(p_peer - p_nccl)))
    assert err <= 1e-6, (cfg, err)      # gradient atomics are unordered: the two runs differ at rounding level
    if rank == 0:
        print('training ok:', cfg, 'max |p2p - nccl| = %.3g' % err, flush=True)
dist.barrier()
dist.destroy_process_group()
if rank == 0:
    print('MULTI_GPU_WORKER_OK', flush=True)
