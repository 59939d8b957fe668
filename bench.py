#!/usr/bin/env python3
"""bench.py -- graph-pairs/s of the Siamese graph-pair similarity hot path on B200.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --steps K --warmup W    # CPU baseline arm (oracle port, nothing of the product loaded)

Headline workload (config.workload): BASELINE.json configs[3] on the configs[0]/[1] model -- all-pairs similarity
ranking, 1 000 query x 10 000 database AIDS-shaped graphs (<= 10 nodes, 29 labels) per GPU, the SimGNN-style chain
(3 GCN + Attention + NTN(16,16) + 16-bin node-node histogram + FC head), top-10 per query.  One "step" = embed every
query and database graph, score every (query, database) pair, select the top-10 ids per query -- what the reference's
test() does with m x n sess.run calls (train.py:57-61).  Nothing is cached across steps.  With N GPUs every rank
holds its own 10 000-graph database shard (weak scaling), queries are replicated, and the per-shard top-10 are
exchanged over peer memory and merged (verified once, untimed, against all-gather + merge: exchange_verified).

Also inside the same JSON line: `default_model` (the same step on the reference's default-flag model), `train` /
`train_simgnn` (configs[2]: data-parallel training, 128 pairs/GPU/step on AIDS10knef-shaped graphs, fwd + bwd +
gradient exchange + Adam, plus a 16 384-pair throughput point), `web` (configs[4]), `metrics` (precision@10 and
Kendall tau kernels on the 1 k x 10 k matrix), `parity` (device scores vs the oracle port on a sample), and a
compact `summary` at the end of the line.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

DEFAULT_LAYERS = [
    'GraphConvolution:output_dim=32,act=relu,dropout=True,bias=True,sparse_inputs=True',
    'GraphConvolution:input_dim=32,output_dim=16,act=identity,dropout=True,bias=True,sparse_inputs=False',
    'Dense:input_dim=16,output_dim=1,dropout=True,act=relu,bias=True',
    'Padding:max_in_dims=10,padding_value=0',
    'NTN:input_dim=10,feature_map_dim=10,inneract=relu,dropout=True,bias=True',
]
TRAIN_LAYERS = DEFAULT_LAYERS[:2] + [  # tuning.py:74-93 (Average pooling: AIDS10k graphs exceed Padding's 10)
    'Average', 'NTN:input_dim=16,feature_map_dim=10,inneract=relu,dropout=True,bias=True']
SIMGNN_LAYERS = [  # BASELINE configs[0]/[1]: SimGNN-style chain (3 GCN + Attention + NTN(16,16) + Histogram(16) + FC head)
    'GraphConvolution:output_dim=64,act=relu,dropout=True,bias=True,sparse_inputs=True',
    'GraphConvolution:input_dim=64,output_dim=32,act=relu,dropout=True,bias=True,sparse_inputs=False',
    'GraphConvolution:input_dim=32,output_dim=16,act=identity,dropout=True,bias=True,sparse_inputs=False',
    'Attention:input_dim=16',
    'NTN:input_dim=16,feature_map_dim=16,inneract=relu,dropout=True,bias=True',
    'Histogram:bins=16',
    'Dense:input_dim=32,output_dim=16,dropout=False,act=relu,bias=True',
    'Dense:input_dim=16,output_dim=8,dropout=False,act=relu,bias=True',
    'Dense:input_dim=8,output_dim=4,dropout=False,act=relu,bias=True',
    'Dense:input_dim=4,output_dim=1,dropout=False,act=identity,bias=True',
]
M_QUERIES, N_DB, TOPK = 1000, 10000, 10
TRAIN_BATCH, TRAIN_GRAPHS = 128, 4096


def peaks():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return d.get('hbm_gbs', 6650.0), 'measured (MEASURED_PEAKS.json)'
    return 6650.0, 'fallback (B200_PROFILING.md)'


class ClockSampler(object):
    """SM clock + throttle reasons sampled through NVML every 5 ms while the timed legs run (the timed
    region of K steps lasts only a few ms, too short for `nvidia-smi -lms`); falls back to nvidia-smi."""
    BITS = {'hw_slowdown': 0x8, 'hw_thermal_slowdown': 0x40, 'sw_thermal_slowdown': 0x20, 'sw_power_cap': 0x4}

    def __init__(self, index):
        self.index, self.rows, self.marks, self.stop_flag, self.thread = index, [], {}, False, None
        self.max_mhz = None

    def _loop_nvml(self, nv, h):
        while not self.stop_flag:
            try:
                sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
                try:
                    rs = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    rs = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                self.rows.append((time.perf_counter(), float(sm), int(rs)))
            except Exception:
                pass
            time.sleep(0.005)

    def _loop_smi(self):
        q = 'clocks.sm,clocks.max.sm,clocks_event_reasons.active'
        while not self.stop_flag:
            try:
                o = subprocess.run(['nvidia-smi', '-i', str(self.index), '--query-gpu=' + q,
                                    '--format=csv,noheader,nounits'], capture_output=True, text=True, timeout=5).stdout
                c = [x.strip() for x in o.strip().split(',')]
                self.max_mhz = float(c[1])
                self.rows.append((time.perf_counter(), float(c[0]), int(c[2], 16)))
            except Exception:
                time.sleep(0.05)

    def start(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            # NVML enumerates physical devices: honour CUDA_VISIBLE_DEVICES when it is a plain index list
            vis = os.environ.get('CUDA_VISIBLE_DEVICES', '')
            idx = self.index
            if vis and all(v.strip().isdigit() for v in vis.split(',')):
                idx = int(vis.split(',')[self.index])
            h = nv.nvmlDeviceGetHandleByIndex(idx)
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            self.thread = threading.Thread(target=self._loop_nvml, args=(nv, h), daemon=True)
        except Exception:
            self.thread = threading.Thread(target=self._loop_smi, daemon=True)
        self.thread.start()

    def mark(self, name):
        self.marks[name] = time.perf_counter()

    def stop(self):
        self.stop_flag = True
        if self.thread:
            self.thread.join(timeout=6)
        t0, t1 = self.marks.get('timed_begin', 0.0), self.marks.get('timed_end', float('inf'))
        l0, l1 = self.marks.get('load_begin', 0.0), self.marks.get('load_end', float('inf'))
        timed = [r for r in self.rows if t0 <= r[0] <= t1]
        load = [r for r in self.rows if l0 <= r[0] <= l1]
        use = load or self.rows
        if not use:
            return {'sm_mhz': None, 'sm_max_mhz': self.max_mhz, 'reasons': ['no clock samples']}
        mask = 0
        for r in use:
            mask |= r[2]
        return {'sm_mhz': float(np.median([r[1] for r in use])), 'sm_max_mhz': self.max_mhz,
                'reasons': sorted(n for n, b in self.BITS.items() if mask & b), 'samples': len(use),
                'samples_in_timed_region': len(timed),
                'window': 'pre-load + timed steps + per-stage timings (GPU continuously busy), 5 ms NVML polling'}


def slice_batch(b, lo, hi):
    """Contiguous sub-batch [lo, hi) of a host PackedBatch (offsets rebased; no per-graph loop)."""
    from graphembedding_b200.native import PackedBatch
    no, rp = np.asarray(b.node_off), np.asarray(b.rowptr)
    n0, n1 = int(no[lo]), int(no[hi])
    e0, e1 = int(rp[n0]), int(rp[n1])
    counts = np.diff(no[lo:hi + 1])
    return PackedBatch((no[lo:hi + 1] - n0).astype(np.int32), np.asarray(b.labels)[n0:n1], (rp[n0:n1 + 1] - e0).astype(np.int32),
                       np.asarray(b.col)[e0:e1], np.asarray(b.val)[e0:e1], int(counts.max()) if hi > lo else 0)


def make_rank_data(rank, world=1, scaling='weak', db_scale=1):
    """-> (encoder, queries, this rank's database shard, id offset of the shard, database graphs over all ranks).
    weak: every rank owns its own N_DB * db_scale graphs; strong (BASELINE configs[3] as written): ONE database of
    N_DB * db_scale graphs split into contiguous shards (distributed.shard_range).  db_scale > 1 tiles the corpus."""
    from graphembedding_b200 import data as D
    from graphembedding_b200 import distributed as DS
    from graphembedding_b200.graphs import NodeFeatureOneHotEncoder, pack_nx_graphs
    from graphembedding_b200.native import PackedBatch
    enc = NodeFeatureOneHotEncoder([], 'type', feat_idx_dic={l: i for i, l in enumerate(D.AIDS_LABELS)})
    q_gs, db_gs = rank_corpora(rank if scaling == 'weak' else 0)
    q, db = pack_nx_graphs(q_gs, enc), pack_nx_graphs(db_gs, enc)
    if db_scale > 1:
        db = PackedBatch.concat([db] * db_scale)
    if scaling == 'weak':
        return enc, q, db, rank * db.n_graphs, world * db.n_graphs
    total = db.n_graphs
    lo, hi = DS.shard_range(total, rank, world)
    return enc, q, slice_batch(db, lo, hi), lo, total


def make_train_data(rank, world, n_batches=8):
    """AIDS10knef-shaped graphs (<= 30 nodes) and n_batches mini-batches of TRAIN_BATCH pairs with
    proxy-GED targets exp(-yeta nGED^2); rank r takes pairs r, r+world, ... of the global stream."""
    from graphembedding_b200 import data as D
    from graphembedding_b200.dist_calculator import proxy_ged
    from graphembedding_b200.graphs import NodeFeatureOneHotEncoder, pack_nx_graphs
    enc = NodeFeatureOneHotEncoder([], 'type', feat_idx_dic={l: i for i, l in enumerate(D.AIDS_LABELS)})
    gs = D.synthetic_corpus(TRAIN_GRAPHS, 2, 30, seed=223, size_skew=False)
    batch = pack_nx_graphs(gs, enc)
    rng = np.random.default_rng(99)
    out = []
    for _ in range(n_batches):
        glob = rng.integers(0, TRAIN_GRAPHS, size=(TRAIN_BATCH * world, 2))
        d = np.array([2.0 * proxy_ged(gs[a], gs[b]) / (gs[a].number_of_nodes() + gs[b].number_of_nodes())
                      for a, b in glob])
        t = np.exp(-0.6 * d * d).astype(np.float32)
        mine = slice(rank, None, world)
        out.append((glob[mine, 0].astype(np.int32), glob[mine, 1].astype(np.int32), t[mine],
                    float(t.astype(np.float64).sum()), float((t.astype(np.float64) ** 2).sum())))
    return enc, batch, out


MODELS = {   # name -> (layer specs, engine keywords, short description)
    'simgnn': (SIMGNN_LAYERS, dict(final_act='sigmoid'),
               'SimGNN-style chain GCN64-GCN32-GCN16-Attention-NTN(16,16)-Histogram(16)-FC(32-16-8-4-1), sigmoid'),
    'default': (DEFAULT_LAYERS, dict(),
                'reference default flags model GCN32-GCN16-Dense1-Pad10-NTN(10,10), sim_kernel'),
}
HEADLINE = 'simgnn'


def oracle_side(gs):
    """networkx graphs -> padded oracle inputs (labels, A_hat, mask, n), with NOTHING of the product loaded: the
    label index and the normalised adjacency come from the oracle's own restatement of graphs.py:55-117."""
    from graphembedding_b200 import data as D          # synthetic generator only (pure Python)
    from oracle import batched as OB
    from oracle import siamese_oracle as O
    idx = {l: i for i, l in enumerate(D.AIDS_LABELS)}
    og = []
    for g in gs:
        pos = {v: i for i, v in enumerate(g.nodes())}
        og.append(O.Graph.from_edges([idx[g.nodes[v]['type']] for v in g.nodes()],
                                     [(pos[u], pos[v]) for u, v in g.edges()]))
    return OB.pad_graphs(og, 29)


def rank_corpora(rank):
    from graphembedding_b200 import data as D
    return (D.synthetic_corpus(M_QUERIES, 2, 10, seed=124, gid_base=10 ** 6),
            D.synthetic_corpus(N_DB, 2, 10, seed=123 + 1000 * rank, gid_base=rank * N_DB))


def cpu_rank_baseline(model, pq, pd, seconds_budget=15.0, threads=None, q_sample=None):
    """Oracle port on the host cores: batched torch-CPU all-pairs ranking step (embedding of the sampled queries
    and of ALL database graphs + scoring + top-k) on a bounded sample of the queries.  pq / pd: padded oracle
    inputs (oracle_side / OB.pad_from_packed).  Returns (pairs/s, cores, sample text)."""
    import torch
    from oracle import batched as OB
    from oracle import siamese_oracle as O
    threads = threads or os.cpu_count()
    torch.set_num_threads(threads)
    spec, kw, _ = MODELS[model]
    layers = O.parse_layer_specs(spec, 29)
    P = O.unflatten(layers, O.init_params(layers, 123), torch.float32)
    head = any(L['kind'] == 'Histogram' for L in layers)
    nq_all, nd_all = int(pq[0].shape[0]), int(pd[0].shape[0])
    qs = nq_all if q_sample is None else min(q_sample, nq_all)
    pq = tuple(t[:qs] for t in pq)
    chunk = 8 if head else 100                     # the [Q, M, Nq, Nd] cell tensor of the histogram must stay small
    t0 = time.perf_counter()
    reps = 0
    while True:
        if head:
            xq, hq = OB.embed(layers, P, *pq, with_nodes=True)
            xd, hd = OB.embed(layers, P, *pd, with_nodes=True)
        else:
            xq, xd = OB.embed(layers, P, *pq), OB.embed(layers, P, *pd)
        for c in range(0, qs, chunk):
            if head:
                sim = OB.score_all_pairs_head(layers, P, xq[c:c + chunk], hq[c:c + chunk], pq[3][c:c + chunk], xd, hd,
                                              pd[3], final_act=kw.get('final_act', 'sim_kernel'))
            else:
                sim = OB.score_all_pairs(layers, P, xq[c:c + chunk], xd)
            torch.topk(sim, TOPK, dim=1)
        reps += 1
        if time.perf_counter() - t0 > seconds_budget or reps >= 50:
            break
    dt = time.perf_counter() - t0
    return qs * nd_all * reps / dt, threads, \
        '{} steps of {} of the {} queries x {} database graphs (embed + score + top-{}), batched torch-CPU fp32 ' \
        'oracle port, inputs pre-padded'.format(reps, qs, nq_all, nd_all, TOPK)


def cpu_per_pair_rate(model, q_gs, db_gs, n_pairs=100):
    """The reference's own control flow: one un-batched forward per (query, database) pair
    (train.py:57-61), single thread."""
    import torch
    from graphembedding_b200 import data as D
    from oracle import siamese_oracle as O
    torch.set_num_threads(1)
    layers = O.parse_layer_specs(MODELS[model][0], 29)
    P = O.unflatten(layers, O.init_params(layers, 123), torch.float32)
    idx = {l: i for i, l in enumerate(D.AIDS_LABELS)}

    def graph(g):
        pos = {v: i for i, v in enumerate(g.nodes())}
        return O.Graph.from_edges([idx[g.nodes[v]['type']] for v in g.nodes()], [(pos[u], pos[v]) for u, v in g.edges()])
    t0 = time.perf_counter()
    for i in range(n_pairs):
        g1, g2 = graph(q_gs[i % len(q_gs)]), graph(db_gs[(7 * i) % len(db_gs)])
        x1, h1 = O.graph_forward(layers, P, g1, 29, torch.float32, with_nodes=True)
        x2, h2 = O.graph_forward(layers, P, g2, 29, torch.float32, with_nodes=True)
        O.pair_forward(layers, P, x1, x2, h1=h1, h2=h2)
    torch.set_num_threads(os.cpu_count())
    return n_pairs / (time.perf_counter() - t0)


def run_reference(args):
    """--impl reference: the reference's CPU path restated (TensorFlow is not installed, so the reference's own
    TF graph cannot run on this box; the oracle port stands in).  Loads nothing of the product: inputs are built
    by the oracle's own preprocessing from the synthetic corpus."""
    rank = int(os.environ.get('RANK', 0))
    if rank != 0:
        return
    q_gs, db_gs = rank_corpora(0)
    pq, pd = oracle_side(q_gs), oracle_side(db_gs)
    vals = []
    for s in range(args.warmup + args.steps):
        v, cores, sample = cpu_rank_baseline(HEADLINE, pq, pd, seconds_budget=5.0, q_sample=64)
        if s >= args.warmup:
            vals.append(v)
    value = float(np.mean(vals))
    pp = cpu_per_pair_rate(HEADLINE, q_gs, db_gs)
    ms = 1000.0 * M_QUERIES * N_DB / value
    print(json.dumps({
        'impl': 'reference', 'metric': 'graph-pairs/sec (all-pairs ranking inference)', 'value': value,
        'unit': 'pairs/s', 'n_gpus': args.gpus, 'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': ms,
        'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
        'config': workload_config(),
        'cpu_baseline': {'value': value, 'unit': 'pairs/s', 'cores': cores, 'kind': 'port', 'sample': sample,
                         'per_pair_reference_structure_pairs_per_s': pp},
        'e2e': {'value': value, 'unit': 'pairs/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}}))


def workload_config(model=None, scaling='weak', db_total=None, world=1):
    model = model or HEADLINE
    cfg = {'workload': 'all-pairs ranking, BASELINE configs[3] on the configs[0]/[1] model: {} queries x {} database '
                       'graphs per GPU, AIDS-shaped (2..10 nodes, 29 labels), {}, top-{}'
                       .format(M_QUERIES, N_DB, MODELS[model][2], TOPK),
           'model': model, 'queries': M_QUERIES, 'db_per_gpu': N_DB, 'topk': TOPK,
           'l2': 'flushed between timed steps (256 MiB write; N > 1: followed by a device-side barrier, both untimed)',
           'ntn_mode': 'reference', 'dropout': 0.0}
    if scaling != 'weak' or (db_total is not None and db_total != N_DB * world):
        cfg['workload'] = cfg['workload'].replace('{} database graphs per GPU'.format(N_DB),
                                                  '{} database graphs in total ({} scaling: contiguous shards)'.format(db_total, scaling)
                                                  if scaling != 'weak' else '{} database graphs per GPU'.format(db_total // world))
        cfg['db_total'] = db_total
        cfg['db_per_gpu'] = db_total // world
    return cfg


def timed_steps(torch, fn, steps, flush, align=None):
    """K device-timed calls of fn with an L2 flush before each; returns total ms.  align (N > 1): a
    device-side barrier between the flush and the start event, so that every rank's timed step starts
    together and the timed region holds the step, not the other ranks' flushes."""
    evs = []
    for _ in range(steps):
        flush.zero_()
        if align is not None:
            align()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        evs.append((a, b))
    torch.cuda.synchronize()
    return sum(a.elapsed_time(b) for a, b in evs)


def peak_tflops():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return d.get('bf16_tflops_sustained', 1400.0), 'measured sustained bf16 (MEASURED_PEAKS.json)'
    return 1400.0, 'fallback (B200_PROFILING.md, sustained)'


def pair_stage_flops(layers, q, db):
    """Algorithmic FLOPs of the pair stage, SURVEY 8(d): per pair NTN 2 D^2 K + 4 D K + 3 K, histogram
    2 N1 N2 d + 2 N1 N2 (summed over the actual node counts), FC head 2 * sum(din * dout).  -> (ntn, hist, head)."""
    pi = [i for i, L in enumerate(layers) if L['kind'] in ('NTN', 'Dot')][0]
    L = layers[pi]
    pairs = float(q.n_graphs) * float(db.n_graphs)
    D, K = (L['input_dim'], L['feature_map_dim']) if L['kind'] == 'NTN' else (16, 1)
    ntn = (2.0 * D * D * K + 4.0 * D * K + 3.0 * K) * pairs if L['kind'] == 'NTN' else 2.0 * D * pairs
    hist = 0.0
    if any(x['kind'] == 'Histogram' for x in layers):
        d = [x for x in layers[:pi] if x['kind'] in ('GraphConvolution', 'Dense')][-1]['output_dim']
        cells = float(np.diff(np.asarray(q.node_off)).sum()) * float(np.diff(np.asarray(db.node_off)).sum())
        hist = (2.0 * d + 2.0) * cells
    head = 2.0 * sum(x['input_dim'] * x['output_dim'] for x in layers[pi + 1:] if x['kind'] == 'Dense') * pairs
    return ntn, hist, head


def rank_bench(model, torch, dist, dev, rank, world, local, args, barrier, max_over_ranks, q, db, flush, clocks=None,
               id_offset=None, n_total=None):
    """The ranking step of one model: device-timed value, per-stage kernel times, roofline of the dominant kernel,
    end-to-end through the host entry points, parity sample against the fp32 oracle port."""
    from graphembedding_b200 import distributed as DS
    from graphembedding_b200.engine import SgnnEngine
    from graphembedding_b200.layers_factory import parse_layers
    spec, kw, desc = MODELS[model]
    n_loc = db.n_graphs
    id_offset = rank * n_loc if id_offset is None else id_offset
    n_total = world * n_loc if n_total is None else n_total
    layers = parse_layers(spec, 29)
    eng = SgnnEngine(layers, 29, device=local, precision=args.precision, **kw)
    eng.init_params(123)
    eng.reserve(max_graphs=max(n_loc, M_QUERIES), max_nodes=max(q.total_nodes, db.total_nodes), max_nnz=db.nnz, max_queries=M_QUERIES,
                max_db=n_loc)
    qd, dbd = q.to(dev), db.to(dev)
    E = eng.emb_dim
    emb_q, emb_db = torch.empty((M_QUERIES, E), device=dev), torch.empty((n_loc, E), device=dev)
    sim = torch.empty((M_QUERIES, n_loc), dtype=torch.float32, device=dev)
    ids_d = torch.empty((M_QUERIES, TOPK), dtype=torch.int32, device=dev)
    vals_d = torch.empty((M_QUERIES, TOPK), dtype=torch.float32, device=dev)
    gids_d, gvals_d = torch.empty_like(ids_d), torch.empty_like(vals_d)   # global top-k after the exchange (N > 1)
    out = {}
    # database sharded over ranks: local top-k, candidate exchange over NVLink peer memory (fallback: NCCL
    # all-gather + merge kernel), global ids
    peer, exchange, verified = None, 'none (1 GPU)', None
    if world > 1:
        try:
            peer, exchange = DS.PeerTopk(eng, M_QUERIES, TOPK, dist), 'p2p push + merge kernels (CUDA IPC peer memory)'
        except Exception as e:   # no peer access between the GPUs: library collective
            sys.stderr.write('bench.py: peer exchange unavailable (%s); using NCCL all-gather\n' % e)
            exchange = 'nccl all-gather + merge kernel'
        flag = torch.tensor([1 if peer is not None else 0], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.item()) == 0 and peer is not None:
            peer.close()
            peer, exchange = None, 'nccl all-gather + merge kernel'

    def step():
        out['ids'], out['vals'] = DS.sharded_rank(eng, qd, dbd, id_offset, TOPK, dist if world > 1 else None,
                                                  emb_q=emb_q, emb_db=emb_db, sim=sim, peer=peer, ids=ids_d, vals=vals_d,
                                                  out_ids=gids_d, out_vals=gvals_d)

    for _ in range(max(3, args.warmup)):
        step()
    torch.cuda.synchronize()
    if world > 1:
        # untimed verification: the peer-memory exchange must give what all-gather + merge kernel gives
        step()
        torch.cuda.synchronize()
        got_i, got_v = out['ids'].clone(), out['vals'].clone()
        li, lv, _ = eng.rank_device(qd, dbd, TOPK, emb_q=emb_q, emb_db=emb_db, sim=sim)
        cv, ci = DS.gather_topk_candidates(lv, li, id_offset, dist)
        wi, wv = eng.topk_merge(cv, ci, TOPK)
        torch.cuda.synchronize()
        ok = torch.tensor([1 if (torch.equal(got_i, wi) and torch.equal(got_v, wv)) else 0], device=dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        verified = bool(int(ok.item()))
        assert verified, 'sharded ranking: exchange result differs from all-gather + merge'
    if clocks is not None:
        clocks.mark('load_begin')
    t_pre = time.perf_counter()           # untimed pre-load: ~0.3 s of back-to-back steps so that the
    while time.perf_counter() - t_pre < 0.3:   # clocks are in their loaded state and get sampled
        for _ in range(20):
            step()
        torch.cuda.synchronize()
    launches0 = eng.launches
    barrier()
    if clocks is not None:
        clocks.mark('timed_begin')
    total_ms = timed_steps(torch, step, args.steps, flush, align=peer.barrier if peer is not None else None)
    barrier()
    if clocks is not None:
        clocks.mark('timed_end')
    launches = eng.launches - launches0
    ms_per_step = max_over_ranks(total_ms) / args.steps
    res = {'model': model, 'value': M_QUERIES * n_total / (ms_per_step * 1e-3), 'unit': 'pairs/s',
           'ms_per_step': ms_per_step, 'gpu_launches': launches, 'exchange': exchange}
    if verified is not None:
        res['exchange_verified'] = verified
    res.update(rank_stages(model, torch, eng, layers, q, db, qd, dbd, emb_q, emb_db, sim, args, flush))
    if clocks is not None:
        clocks.mark('load_end')
    res['e2e'] = rank_e2e(torch, eng, q, db, dev, world, args, barrier, max_over_ranks, out.get('ids'), n_total)
    if world == 1:
        res['parity'] = rank_parity(model, torch, eng, layers, q, db, qd, dbd)
    if peer is not None:
        peer.close()
    return res


def committed_traffic(kernel):
    """dram bytes per launch of `kernel` from the committed ncu capture of this round (profiles/traffic.json, written
    from the --set full reports next to it), or None."""
    p = os.path.join(ROOT, 'profiles', 'traffic.json')
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f).get(kernel)
    return None


def rank_stages(model, torch, eng, layers, q, db, qd, dbd, emb_q, emb_db, sim, args, flush):
    """Per-stage kernel times of the ranking step (each timed alone on the launching stream, L2 flushed) and the
    roofline of the dominant kernel."""
    K = args.steps
    ntn_f, hist_f, head_f = pair_stage_flops(layers, q, db)
    t_ms = timed_steps(torch, lambda: eng.topk(sim, TOPK), K, flush) / K
    if eng.has_head:
        eq, rq = eng.embed_nodes(qd)
        ed, rd = eng.embed_nodes(dbd)
        e_ms = timed_steps(torch, lambda: (eng.embed_nodes(qd), eng.embed_nodes(dbd)), K, flush) / K
        for _ in range(3):
            eng.score_all_pairs_nodes(eq, rq, qd, ed, rd, dbd, out=sim)
        s_ms = timed_steps(torch, lambda: eng.score_all_pairs_stages(eq, ed, sim, eng.STAGE_SCORE), K, flush) / K
        p_ms = timed_steps(torch, lambda: eng.score_all_pairs_stages(eq, ed, sim, eng.STAGE_PREP_QUERIES |
                                                                     eng.STAGE_PREP_DATABASE), K, flush) / K
        stage = {'embed_q_plus_db_with_node_rows': e_ms, 'operand_prep_q_plus_db': p_ms,
                 'scorer_stage_histogram_plus_head': s_ms, 'topk': t_ms}
        peak, peak_src = peak_tflops()
        tc_hist = args.precision != 'fp32'
        if tc_hist:
            cnt = torch.zeros((q.n_graphs, db.n_graphs, 16), dtype=torch.uint8, device=sim.device)
            eng.hist_all_pairs(rq, qd, rd, dbd, out=cnt, stages=3)
            k_ms = timed_steps(torch, lambda: eng.hist_all_pairs(rq, qd, rd, dbd, out=cnt, stages=4), K, flush) / K
            stage['hist_tc_kernel'] = k_ms
            stage['head_kernel_by_difference'] = s_ms - k_ms
            kernel, k_flops = 'hist_tc_kernel (tcgen05 node-node GEMM + sigmoid/bin epilogue)', hist_f
            del cnt
        else:
            k_ms, kernel, k_flops = s_ms, 'allpairs_head_kernel (fp32 CUDA cores)', ntn_f + hist_f + head_f
        achieved = k_flops / (k_ms * 1e-3) / 1e12
        # what actually bounds the binning epilogue: 1.5 MUFU (ex2 + half a shared rcp) per node-node cell on the
        # XU pipe (16 lanes / clk / SM)
        cells = hist_f / (2.0 * eng.node_dim + 2.0)
        sm_hz = 148 * 1.9e9
        roofline = {'bound': 'tensor', 'achieved': achieved, 'peak': peak, 'unit': 'TFLOP/s', 'frac': achieved / peak,
                    'traffic': committed_traffic('hist_tc_kernel' if tc_hist else 'allpairs_head_kernel'),
                    'kernel': kernel, 'kernel_ms': k_ms, 'peak_source': peak_src,
                    'algorithmic_flops_per_launch': k_flops,
                    'pair_stage': {'algorithmic_gflop': (ntn_f + hist_f + head_f) / 1e9, 'ms': s_ms,
                                   'achieved_tflops': (ntn_f + hist_f + head_f) / (s_ms * 1e-3) / 1e12,
                                   'frac': (ntn_f + hist_f + head_f) / (s_ms * 1e-3) / 1e12 / peak},
                    'epilogue_floor': {'node_node_cells': cells, 'mufu_per_cell': 1.5,
                                       'xu_floor_ms_at_1.9GHz': cells * 1.5 / 16.0 / sm_hz * 1e3,
                                       'note': 'the kernel is bound by the sigmoid/bin epilogue (XU + issue), not by the MMA'},
                    'stage_ms': stage}
        return {'roofline': roofline}
    for _ in range(3):
        eng.score_all_pairs(emb_q, emb_db, out=sim)
    # the scorer kernel proper (operands prepared above): the launch the roofline is quoted on
    k_ms = timed_steps(torch, lambda: eng.score_all_pairs_stages(emb_q, emb_db, sim, eng.STAGE_SCORE), K, flush) / K
    p_ms = timed_steps(torch, lambda: eng.score_all_pairs_stages(emb_q, emb_db, sim, eng.STAGE_PREP_QUERIES |
                                                                 eng.STAGE_PREP_DATABASE), K, flush) / K
    e_ms = timed_steps(torch, lambda: (eng.embed(qd, out=emb_q), eng.embed(dbd, out=emb_db)), K, flush) / K
    peak, peak_src = peaks()
    alg_bytes = 4.0 * M_QUERIES * db.n_graphs + 4.0 * (M_QUERIES + db.n_graphs) * eng.emb_dim
    achieved = alg_bytes / (k_ms * 1e-3) / 1e9
    tc = args.precision != 'fp32'
    return {'roofline': {'bound': 'hbm', 'achieved': achieved, 'peak': peak, 'unit': 'GB/s', 'frac': achieved / peak,
                         'traffic': committed_traffic('allpairs_tc_kernel' if tc else 'allpairs_fp32_kernel'),
                         'kernel': 'allpairs_tc_kernel (tcgen05 scorer, SGNN_STAGE_SCORE)' if tc else 'allpairs_fp32_kernel',
                         'kernel_ms': k_ms, 'peak_source': peak_src, 'algorithmic_bytes_per_launch': alg_bytes,
                         'stage_ms': {'embed_q_plus_db': e_ms, 'operand_prep_q_plus_db': p_ms, 'scorer': k_ms, 'topk': t_ms}}}


def rank_e2e(torch, eng, q, db, dev, world, args, barrier, max_over_ranks, dev_ids, n_total):
    """The same step through the host C-ABI entry points: pinned HOST buffers in, top-k out, wall clock."""
    def timed(fn):
        t_w, n_w = time.perf_counter(), 0
        while n_w < args.warmup or time.perf_counter() - t_w < 0.1:   # warm-up: >= W calls and >= 0.1 s (clocks settle)
            fn()
            n_w += 1
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            fn()
        barrier()
        return max_over_ranks((time.perf_counter() - t0) * 1e3) / args.steps
    qp, dbp = q.pin(), db.pin()
    ids_h = torch.empty((M_QUERIES, TOPK), dtype=torch.int32).pin_memory()
    vals_h = torch.empty((M_QUERIES, TOPK), dtype=torch.float32).pin_memory()
    e2e_ms = timed(lambda: eng.rank_host(qp, dbp, TOPK, ids=ids_h, vals=vals_h))
    if world == 1 and dev_ids is not None:
        assert torch.equal(ids_h.to(dev), dev_ids), 'host entry point and device path disagree'
    d2h = 2 * 4 * M_QUERIES * TOPK
    csr = {'value': M_QUERIES * n_total / (e2e_ms * 1e-3), 'unit': 'pairs/s', 'ms_per_step': e2e_ms,
           'h2d_bytes_per_step': q.nbytes() + db.nbytes(), 'd2h_bytes_per_step': d2h,
           'entry_point': 'sgnn_rank_host (int32 / float CSR arrays from pinned host memory)'}
    # the same step with the batches in the compact wire format (sgnn_compact_batch: byte-wide fields + value
    # dictionary, built once per dataset on the host like the CSR itself; expanded on the device every step)
    qc, dbc = q.compact(), db.compact()
    e2e = csr
    if qc is not None and dbc is not None:
        both = torch.empty((2, M_QUERIES, TOPK), dtype=torch.int32).pin_memory()   # ids | values back to back: one D2H copy
        ids_c, vals_c = both[0], both[1].view(torch.float32)
        c_ms = timed(lambda: eng.rank_host_compact(qc, dbc, TOPK, ids=ids_c, vals=vals_c))
        assert torch.equal(ids_c, ids_h), 'compact and plain host entry points disagree'
        e2e = {'value': M_QUERIES * n_total / (c_ms * 1e-3), 'unit': 'pairs/s', 'ms_per_step': c_ms,
               'h2d_bytes_per_step': int(qc.numel() + dbc.numel()), 'd2h_bytes_per_step': d2h,
               'entry_point': 'sgnn_rank_host_compact (compact wire blobs from pinned host memory, expanded on the device)',
               'plain_csr': csr}
    if world > 1:
        e2e['note'] = 'per-rank host entry point on its own shard (host in, local top-k out); the cross-rank merge is in value'
    return e2e


def metrics_bench(torch, dev, args, flush):
    """precision@10 and Kendall tau of the 1 k x 10 k similarity matrix on the device (src/metrics.py:14-35,
    build-defined tau-b): both read two m x n fp32 matrices once.  precision@10 goes through the small-k form
    (two sgnn_topk launches + sgnn_precision_from_topk); the all-k sorting kernel is timed beside it."""
    from graphembedding_b200.engine import SgnnEngine
    from graphembedding_b200.layers_factory import parse_layers
    eng = SgnnEngine(parse_layers(DEFAULT_LAYERS, 29), 29, device=dev.index)
    g = torch.Generator(device=dev).manual_seed(1)
    true_sim = torch.rand((M_QUERIES, N_DB), device=dev, generator=g)
    pred_sim = (true_sim + 0.05 * torch.randn((M_QUERIES, N_DB), device=dev, generator=g)).clamp_(0, 1)
    ks = np.array([TOPK], np.int32)
    K = max(3, args.steps // 4)
    for _ in range(2):
        eng.precision_at_ks(true_sim, pred_sim, ks)
        eng.kendall_tau(true_sim, pred_sim)
    p_ms = timed_steps(torch, lambda: eng.precision_at_ks(true_sim, pred_sim, ks), K, flush) / K
    ks_all = np.array([TOPK, N_DB - 1], np.int32)          # a k beyond 64: the sorting kernel (every k in one pass)
    eng.precision_at_ks(true_sim, pred_sim, ks_all)
    pa_ms = timed_steps(torch, lambda: eng.precision_at_ks(true_sim, pred_sim, ks_all), K, flush) / K
    k_ms = timed_steps(torch, lambda: eng.kendall_tau(true_sim, pred_sim), K, flush) / K
    peak, _ = peaks()
    by = 2.0 * 4.0 * M_QUERIES * N_DB
    return {'precision_at_10_ms': p_ms, 'precision_all_k_sorting_kernel_ms': pa_ms, 'kendall_tau_ms': k_ms, 'algorithmic_bytes': by,
            'precision_at_10_hbm_frac': by / (p_ms * 1e-3) / 1e9 / peak, 'kendall_tau_hbm_frac': by / (k_ms * 1e-3) / 1e9 / peak}


def run_cuda(args):
    import torch
    import torch.distributed as dist
    world = int(os.environ.get('WORLD_SIZE', 1))
    rank = int(os.environ.get('RANK', 0))
    local = int(os.environ.get('LOCAL_RANK', 0))
    if not torch.cuda.is_available():
        raise SystemExit('bench.py: no CUDA device (the product path has no CPU fallback)')
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    # a real (non-legacy) stream: sgnn_rank_device replays its step as a CUDA graph, and stream
    # capture is not allowed on the legacy default stream
    torch.cuda.set_stream(torch.cuda.Stream())
    # rank 0 prints ONE JSON line on stdout: libraries that write to file descriptor 1 (NCCL's version banner)
    # are sent to stderr for the duration of the run, the JSON line goes to the saved descriptor
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    enc, q, db, id_offset, n_total = make_rank_data(rank, world, args.scaling, args.db_scale)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    clocks = ClockSampler(local)
    clocks.start()
    common = (torch, dist, dev, rank, world, local, args, barrier, max_over_ranks)
    head = rank_bench(HEADLINE, *common, q, db, flush, clocks=clocks, id_offset=id_offset, n_total=n_total)
    clk = clocks.stop()
    extra = {}
    if not args.headline_only:
        other = rank_bench('default', *common, q, db, flush, id_offset=id_offset, n_total=n_total)
        extra['default_model'] = {k: other[k] for k in ('value', 'unit', 'ms_per_step', 'gpu_launches', 'roofline', 'e2e', 'parity')
                                  if k in other}
        extra['default_model']['workload'] = MODELS['default'][2]
        extra['train'] = run_train('avg_ntn', *common)
        extra['train_simgnn'] = run_train('simgnn', *common)
        extra['web'] = run_web(torch, dev, rank, world, local, args, barrier, max_over_ranks, flush)
        if world == 1:
            extra['metrics'] = metrics_bench(torch, dev, args, flush)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return
    cfg = workload_config(scaling=args.scaling, db_total=n_total, world=world)
    line = {'metric': 'graph-pairs/sec (all-pairs ranking inference)', 'value': head['value'], 'unit': 'pairs/s',
            'n_gpus': world, 'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': head['ms_per_step'],
            'higher_is_better': True, 'scaling': args.scaling, 'vs_baseline': None,
            'dtype': 'f32' if args.precision == 'fp32' else args.precision, 'data': 'synthetic',
            'config': cfg, 'clocks': clk, 'e2e': head['e2e'], 'gpu_launches': head['gpu_launches'],
            'roofline': head['roofline'], 'exchange': head['exchange']}
    for k in ('exchange_verified', 'parity'):
        if k in head:
            line[k] = head[k]
    if world == 1 and not args.no_cpu_baseline:
        from oracle import batched as OB
        v, cores, sample = cpu_rank_baseline(HEADLINE, OB.pad_from_packed(q),
                                             OB.pad_from_packed(slice_batch(db, 0, min(db.n_graphs, N_DB))), q_sample=64)
        q_gs, db_gs = rank_corpora(0)
        line['cpu_baseline'] = {'value': v, 'unit': 'pairs/s', 'cores': cores, 'kind': 'port', 'sample': sample,
                                'per_pair_reference_structure_pairs_per_s': cpu_per_pair_rate(HEADLINE, q_gs, db_gs)}
    line.update(extra)
    # the driver keeps the last ~1 500 characters of the line: the numbers of every sub-benchmark once more, compactly
    summ = {'headline_pairs_per_s': head['value'], 'headline_ms': head['ms_per_step'], 'e2e_pairs_per_s': head['e2e']['value'],
            'roofline_frac': head['roofline']['frac'], 'exchange': head['exchange']}
    if 'exchange_verified' in head:
        summ['exchange_verified'] = head['exchange_verified']
    for k in ('default_model', 'train', 'train_simgnn', 'web'):
        if k in extra:
            summ[k + '_pairs_per_s'] = extra[k]['value']
            summ[k + '_ms'] = extra[k]['ms_per_step']
    line['summary'] = summ
    sys.stdout.flush()
    os.write(json_fd, (json.dumps(line) + '\n').encode())


def run_web(torch, dev, rank, world, local, args, barrier, max_over_ranks, flush):
    """BASELINE configs[4]: Web-shaped larger graphs (N ~ U[100, 500], E ~ 2N, 32 labels) stressing the CSR
    gather (large-graph embedding kernel, node rows in a global workspace) and the node-node similarity
    histogram.  One step = embed WEB_GRAPHS graphs (pooled embeddings + node rows), NTN-score WEB_PAIRS listed
    pairs and build their 16-bin histograms."""
    from graphembedding_b200 import data as D
    from graphembedding_b200.engine import SgnnEngine
    from graphembedding_b200.graphs import NodeFeatureOneHotEncoder, pack_nx_graphs
    from graphembedding_b200.layers_factory import parse_layers
    WEB_GRAPHS, WEB_PAIRS = 256, 4096
    W = type('WebBench', (D.WebData,), {'num_train': WEB_GRAPHS})
    gs = W(True).graphs
    enc = NodeFeatureOneHotEncoder(gs, 'type', feat_idx_dic={l: i for i, l in enumerate(D.WebData.labels)})
    batch = pack_nx_graphs(gs, enc)
    eng = SgnnEngine(parse_layers(TRAIN_LAYERS, 32), 32, device=local, precision='bf16x3')   # histogram on the tensor cores
    eng32 = SgnnEngine(parse_layers(TRAIN_LAYERS, 32), 32, device=local)                      # ... and the CUDA-core kernel beside it
    for e in (eng, eng32):
        e.init_params(123)
        e.reserve(max_graphs=WEB_GRAPHS, max_nodes=batch.total_nodes, max_nnz=batch.nnz, max_pairs=WEB_PAIRS)
    bd = batch.to(dev)
    rng = np.random.default_rng(7 + rank)
    g1 = torch.as_tensor(rng.integers(0, WEB_GRAPHS, WEB_PAIRS).astype(np.int32), device=dev)
    g2 = torch.as_tensor(rng.integers(0, WEB_GRAPHS, WEB_PAIRS).astype(np.int32), device=dev)
    out = {}

    def step():
        emb, rows = eng.embed_nodes(bd)
        out['raw'], out['sim'] = eng.score_pairs(emb, g1, g2)
        out['hist'], _ = eng.pair_histogram(rows, bd.node_off, g1, g2, bins=16)

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    barrier()
    ms = max_over_ranks(timed_steps(torch, step, args.steps, flush)) / args.steps
    e_ms = timed_steps(torch, lambda: eng.embed_nodes(bd), args.steps, flush) / args.steps
    emb, rows = eng.embed_nodes(bd)
    h_ms = timed_steps(torch, lambda: eng.pair_histogram(rows, bd.node_off, g1, g2, bins=16), args.steps, flush) / args.steps
    _, c_tc = eng.pair_histogram(rows, bd.node_off, g1, g2, bins=16)
    _, c_32 = eng32.pair_histogram(rows, bd.node_off, g1, g2, bins=16)
    h32_ms = timed_steps(torch, lambda: eng32.pair_histogram(rows, bd.node_off, g1, g2, bins=16), args.steps, flush) / args.steps
    moved = int((c_tc - c_32).abs().sum().item()) // 2     # cells that the two kernels put into different bins (bin-edge band)
    tpeak, tpeak_src = peak_tflops()
    # listed-pair TRAINING on the same graphs (the arms' buffers of the per-pair kernel in a global workspace): 128 pairs
    TB = 128
    tt = torch.as_tensor(np.random.default_rng(3).uniform(0.05, 1.0, TB).astype(np.float32), device=dev)
    tl = torch.zeros(4, device=dev)

    def tstep():
        tl.zero_()
        eng32.train_step(bd, g1[:TB], g2[:TB], tt, TB, 1.0, 1.0, loss=tl)
        eng32.apply_adam(1.0, tl)
    for _ in range(3):
        tstep()
    t_ms = timed_steps(torch, tstep, max(3, args.steps // 2), flush) / max(3, args.steps // 2)
    no = batch.node_off
    n1 = np.diff(no)[g1.cpu().numpy()].astype(np.float64)
    n2 = np.diff(no)[g2.cpu().numpy()].astype(np.float64)
    cells = float((n1 * n2).sum())
    peak, _ = peaks()
    gather_bytes = float(batch.algorithmic_bytes() + 4 * batch.total_nodes * 16 + 4 * WEB_GRAPHS * 16)
    return {'workload': 'BASELINE configs[4]: {} Web-shaped graphs (N~U[100,500], mean {:.0f} nodes, {} non-zeros), '
                        'GCN32-GCN16-Average-NTN(16,10) + 16-bin node-node histogram on {} listed pairs per GPU'
                        .format(WEB_GRAPHS, batch.total_nodes / WEB_GRAPHS, batch.nnz, WEB_PAIRS),
            'value': WEB_PAIRS * world / (ms * 1e-3), 'unit': 'pairs/s', 'ms_per_step': ms,
            'graphs_per_s': WEB_GRAPHS * world / (ms * 1e-3),
            'stage_ms': {'embed_large_with_node_rows': e_ms, 'pair_histogram': h_ms, 'pair_histogram_cuda_core': h32_ms},
            'roofline': {'bound': 'tensor', 'kernel': 'hist_pairs_tc_kernel (+ tile prep, 2 scans, finalize)',
                         'achieved': 2.0 * 16 * cells / (h_ms * 1e-3) / 1e12, 'peak': tpeak, 'unit': 'TFLOP/s',
                         'frac': 2.0 * 16 * cells / (h_ms * 1e-3) / 1e12 / tpeak, 'traffic': None, 'peak_source': tpeak_src,
                         'note': 'algorithmic FLOPs 2 d per node-node cell (SURVEY 8(d)); the kernel is bound by the '
                                 'sigmoid / bin epilogue (7.75 instructions per cell), not by the six bf16 products'},
            'parity': {'cells_binned_differently_than_cuda_core_kernel': moved, 'cells': cells},
            'train': {'pairs_per_step': TB, 'ms_per_step': t_ms, 'value': TB * world / (t_ms * 1e-3), 'unit': 'pairs/s',
                      'note': 'fwd + bwd + Adam on 128 listed pairs of the Web-shaped graphs (train_team_kernel, arm '
                              'buffers in a global workspace), fp32'},
            'embed_gather': {'algorithmic_bytes': gather_bytes, 'achieved_GBps': gather_bytes / (e_ms * 1e-3) / 1e9,
                             'hbm_peak_GBps': peak},
            'histogram': {'node_node_cells': cells, 'cells_per_s': cells / (h_ms * 1e-3),
                          'gflops_fp32': 2.0 * 16 * cells / (h_ms * 1e-3) / 1e9}}


TRAIN_MODELS = {'avg_ntn': (TRAIN_LAYERS, 'GCN32-GCN16-Average-NTN(16,10)'),
                'simgnn': (SIMGNN_LAYERS, 'GCN64-GCN32-GCN16-Attention-NTN(16,16)-Histogram(16)-FC(32-16-8-4-1)')}


def train_cpu_baseline(name, batch, batches, seconds_budget=8.0):
    """Oracle port of one optimiser step on the host cores: batched torch-CPU autograd + TF-1 Adam
    (oracle/batched.py train_step) for chains without a head, the per-pair fp64 restatement for the SimGNN-style head."""
    import torch
    from oracle import batched as OB
    from oracle import siamese_oracle as O
    layers = O.parse_layer_specs(TRAIN_MODELS[name][0], 29)
    flat = O.init_params(layers, 123)
    g1, g2, t = batches[0][0], batches[0][1], batches[0][2]
    if name == 'avg_ntn':
        torch.set_num_threads(os.cpu_count())
        lab, A, mask, n = OB.pad_from_packed(batch)
        adam = O.TFAdam(len(flat))
        t0, reps = time.perf_counter(), 0
        while time.perf_counter() - t0 < seconds_budget and reps < 200:
            flat, _ = OB.train_step(layers, flat, adam, lab, A, mask, n, torch.as_tensor(g1.astype(np.int64)),
                                    torch.as_tensor(g2.astype(np.int64)), torch.as_tensor(t), loss_mode='paired')
            reps += 1
        dt = time.perf_counter() - t0
        return {'value': reps * len(g1) / dt, 'unit': 'pairs/s', 'cores': os.cpu_count(), 'kind': 'port',
                'sample': '{} optimiser steps of {} pairs, batched torch-CPU autograd + TF-1 Adam over all {} graphs'
                          .format(reps, len(g1), batch.n_graphs)}
    torch.set_num_threads(1)
    no, rp = np.asarray(batch.node_off), np.asarray(batch.rowptr)

    def graph(g):
        lo, hi = no[g], no[g + 1]
        r = rp[lo:hi + 1]
        rows = np.repeat(np.arange(hi - lo), np.diff(r))
        return O.Graph(np.asarray(batch.labels)[lo:hi], rows, np.asarray(batch.col)[r[0]:r[-1]], np.asarray(batch.val)[r[0]:r[-1]])
    k = 16
    a1, a2 = [graph(g) for g in g1[:k]], [graph(g) for g in g2[:k]]
    t0 = time.perf_counter()
    O.loss_and_grads(layers, flat, a1, a2, t[:k], 29, final_act='sigmoid', weight_decay=0.0, loss_mode='paired')
    dt = time.perf_counter() - t0
    torch.set_num_threads(os.cpu_count())
    return {'value': k / dt, 'unit': 'pairs/s', 'cores': 1, 'kind': 'port',
            'sample': 'forward + fp64 autograd backward of {} pairs, per-pair restatement (the reference control flow)'.format(k)}


def run_train(name, torch, dist, dev, rank, world, local, args, barrier, max_over_ranks, cpu=True):
    """BASELINE configs[2]: data-parallel training, TRAIN_BATCH pairs per GPU per step (fwd + bwd + gradient
    exchange + Adam), plus a throughput point at BIG pairs per step that fills the GPU."""
    from graphembedding_b200 import distributed as DS
    from graphembedding_b200.engine import SgnnEngine
    from graphembedding_b200.layers_factory import parse_layers
    spec, desc = TRAIN_MODELS[name]
    enc, batch, batches = make_train_data(rank, world)
    kw = dict(final_act='sigmoid') if name == 'simgnn' else {}
    eng = SgnnEngine(parse_layers(spec, 29), 29, device=local, loss_mode='paired', ntn_mode='reference', **kw)
    eng.init_params(123)
    bd = batch.to(dev)
    feeds = [(torch.as_tensor(a, device=dev), torch.as_tensor(b, device=dev), torch.as_tensor(t, device=dev), s1, s2)
             for a, b, t, s1, s2 in batches]
    loss = torch.zeros(4, dtype=torch.float32, device=dev)
    grads = eng.device_buffers()[1]
    gb = TRAIN_BATCH * world
    peer, allreduce = None, 'none (1 GPU)'
    if world > 1:   # gradient push over peer memory fused into Adam; fallback: NCCL all-reduce + Adam
        try:
            peer, allreduce = DS.PeerTopk(eng, 0, 0, dist), 'p2p push fused into Adam (CUDA IPC peer memory)'
        except Exception as e:
            sys.stderr.write('bench.py: peer exchange unavailable (%s); using NCCL all-reduce\n' % e)
        flag = torch.tensor([1 if peer is not None else 0], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.item()) == 0:
            if peer is not None:
                peer.close()
            peer, allreduce = None, 'nccl all-reduce'

    def step(i):
        g1, g2, t, s1, s2 = feeds[i % len(feeds)]
        loss.zero_()
        DS.data_parallel_step(eng, bd, g1, g2, t, gb, s1, s2, loss, dist if world > 1 else None, grads=grads,
                              peer=peer)

    def timed(fn, n):
        for i in range(10):
            fn(i)
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for i in range(n):
            fn(i)
        b.record()
        barrier()
        return max_over_ranks(a.elapsed_time(b)) / n

    n = max(args.steps * 10, 100)
    l0 = eng.launches
    ms = timed(step, n)
    launches = (eng.launches - l0) * n // (n + 10)
    final_loss = float(loss[0] + loss[1])
    # algorithmic bytes per pair: both graphs' CSR + 4 B target + 4 B output
    no, rp = np.asarray(batch.node_off), np.asarray(batch.rowptr)
    gbytes = 4.0 * np.diff(no) + 4.0 * (np.diff(no) + 1) + 8.0 * (rp[no[1:]] - rp[no[:-1]])

    def roof(pairs_g1, pairs_g2, step_ms):
        by = float(gbytes[pairs_g1].sum() + gbytes[pairs_g2].sum() + 8.0 * len(pairs_g1))
        peak, src = peaks()
        ach = by / (step_ms * 1e-3) / 1e9
        return {'bound': 'hbm', 'achieved': ach, 'peak': peak, 'unit': 'GB/s', 'frac': ach / peak, 'traffic': None,
                'kernel': 'train_team_kernel + adam_kernel (one optimiser step)', 'kernel_ms': step_ms,
                'algorithmic_bytes_per_step': by, 'peak_source': src}
    res = {'allreduce': allreduce, 'model': desc, 'value': TRAIN_BATCH * world / (ms * 1e-3), 'unit': 'pairs/s',
           'ms_per_step': ms, 'steps': n, 'gpu_launches': launches, 'final_loss': final_loss, 'dtype': 'f32',
           'roofline': roof(batches[0][0], batches[0][1], ms)}
    if world == 1:
        # the same optimiser steps as ONE CUDA graph of len(feeds) steps (engine.capture_train_steps: Adam's step count on
        # the device): the eager 128-pair step above is bound by the launch path (three launches per step)
        gfeeds = [(f[0], f[1], f[2], gb, f[3], f[4]) for f in feeds]
        graph = eng.capture_train_steps(bd, gfeeds, loss)
        reps = max(2, n // len(gfeeds))
        gms = timed(lambda i: graph.replay(), reps) / len(gfeeds)
        eng.adam_device_step(False)
        res['eager_ms_per_step'] = ms
        res['ms_per_step'] = gms
        res['value'] = TRAIN_BATCH / (gms * 1e-3)
        res['graph'] = '{} optimiser steps per CUDA-graph launch'.format(len(gfeeds))
        res['roofline'] = roof(batches[0][0], batches[0][1], gms)
    # ---- throughput point: BIG pairs per step (random pairs of the same graphs, random targets), no exchange ----
    BIG = 16384
    rng = np.random.default_rng(5 + rank)
    b1, b2 = rng.integers(0, TRAIN_GRAPHS, BIG).astype(np.int32), rng.integers(0, TRAIN_GRAPHS, BIG).astype(np.int32)
    bt = rng.uniform(0.05, 1.0, BIG).astype(np.float32)
    d1, d2, dt_ = torch.as_tensor(b1, device=dev), torch.as_tensor(b2, device=dev), torch.as_tensor(bt, device=dev)
    ts, ts2 = float(bt.astype(np.float64).sum()), float((bt.astype(np.float64) ** 2).sum())

    def big(i):
        loss.zero_()
        eng.train_step(bd, d1, d2, dt_, BIG, ts, ts2, loss=loss)
        eng.apply_adam(1.0, loss)
    bms = timed(big, max(args.steps, 20))
    res['large_batch'] = {'pairs_per_step': BIG, 'value': BIG * world / (bms * 1e-3), 'unit': 'pairs/s', 'ms_per_step': bms,
                          'roofline': roof(b1, b2, bms), 'note': 'per-GPU replicas, no gradient exchange'}
    if cpu and world == 1 and not args.no_cpu_baseline:
        res['cpu_baseline'] = train_cpu_baseline(name, batch, batches)
    if peer is not None:
        peer.close()
    return res


def rank_parity(model, torch, eng, layers, q, db, qd, dbd, nq=64, nd=1000):
    """Device scores of a sample (first nq queries x first nd database graphs) against the fp32 oracle port."""
    from oracle import batched as OB
    from oracle import siamese_oracle as O
    from scipy.stats import kendalltau
    spec, kw, _ = MODELS[model]
    ol = O.parse_layer_specs(spec, 29)
    P = O.unflatten(ol, O.init_params(ol, 123), torch.float32)
    flat = O.init_params(ol, 123)
    eng2_params = eng.get_params()
    eng.set_params(flat)                       # identical weights on both sides
    ids, vals, sim = eng.rank_device(qd, dbd, TOPK)
    got = sim[:nq, :nd].cpu().numpy().astype(np.float64)
    eng.set_params(eng2_params)
    pq = tuple(t[:nq] for t in OB.pad_from_packed(q))
    pd = OB.pad_from_packed(slice_batch(db, 0, min(nd, db.n_graphs)))
    if eng.has_head:
        xq, hq = OB.embed(ol, P, *pq, with_nodes=True)
        xd, hd = OB.embed(ol, P, *pd, with_nodes=True)
        ref = OB.score_all_pairs_head(ol, P, xq, hq, pq[3], xd, hd, pd[3], final_act=kw.get('final_act', 'sim_kernel'))
    else:
        ref = OB.score_all_pairs(ol, P, OB.embed(ol, P, *pq), OB.embed(ol, P, *pd))
    ref = ref.numpy().astype(np.float64)
    err = np.abs(got - ref)
    top_g = np.argsort(-got, axis=1, kind='stable')[:, :TOPK]
    top_r = np.argsort(-ref, axis=1, kind='stable')[:, :TOPK]
    prec = float(np.mean([len(set(a) & set(b)) / float(TOPK) for a, b in zip(top_g, top_r)]))
    tau = float(np.mean([kendalltau(a, b).statistic for a, b in zip(got, ref)]))
    return {'sample': '{} queries x {} database graphs vs the fp32 torch-CPU oracle port'.format(nq, nd),
            'max_abs_err': float(err.max()), 'max_rel_err': float((err / np.maximum(np.abs(ref), 1e-6)).max()),
            'mse': float((err ** 2).mean()), 'frac_cells_err_gt_1e-4': float((err > 1e-4).mean()),
            'prec_at_10_vs_oracle': prec, 'kendall_tau_vs_oracle': tau,
            'note': 'cells above 1e-4 are pairs with a node-node score on a histogram bin edge (one count moves)'}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='cuda', choices=['cuda', 'reference'])
    ap.add_argument('--scaling', default='weak', choices=['weak', 'strong'],
                    help='weak: N_DB database graphs PER GPU (default); strong: one database split into contiguous shards')
    ap.add_argument('--db-scale', type=int, default=1, help='tile the 10 000-graph database corpus this many times')
    ap.add_argument('--headline-only', action='store_true', help='skip the sub-benchmarks (default model, train, web, metrics)')
    ap.add_argument('--precision', default='bf16x3', choices=['fp32', 'bf16x3', 'bf16'],
                    help='all-pairs scorer: bf16x3 = tcgen05 with split-bf16 operands and fp32 accumulation (fp32-grade, '
                         'gated at 2e-4 rel against the fp32 CUDA-core scorer), fp32 = CUDA-core parity mode')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    # a hung kernel must not eat the GPU lease: dump every thread's stack and exit after the watchdog
    import faulthandler
    faulthandler.dump_traceback_later(int(os.environ.get('BENCH_WATCHDOG_S', '900')), exit=True)
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_cuda(args)


if __name__ == '__main__':
    main()
