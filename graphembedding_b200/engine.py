"""Thin object wrapper over the libsgnn handle.  PyTorch is used for device memory and
streams only; every computation is a C-ABI call into the hand-written CUDA kernels."""
import ctypes as C

import numpy as np

from . import native
from .native import PackedBatch, check, lib, _ptr


def _stream_ptr(stream):
    if stream is None:
        import torch
        return torch.cuda.current_stream().cuda_stream
    return stream if isinstance(stream, int) else stream.cuda_stream


class SgnnEngine(object):
    def __init__(self, layers, input_dim, **kw):
        import torch
        if not torch.cuda.is_available():
            raise native.SgnnError('no CUDA device: the sm_100a path has no CPU fallback')
        self.layers = layers
        pair_at = [i for i, L in enumerate(layers) if L['kind'] in ('NTN', 'Dot')]
        self.has_head = bool(pair_at) and pair_at[0] < len(layers) - 1   # Histogram / Dense after the NTN
        self.cfg = native.make_config(layers, input_dim, **kw)
        self.device = torch.device('cuda', self.cfg.device)
        h = C.c_void_p()
        rc = lib().sgnn_create(C.byref(self.cfg), C.byref(h))
        if rc != 0:
            raise native.SgnnError(lib().sgnn_last_error(None).decode())
        self.h = h
        self.precision = 'fp32' if self.cfg.precision == 0 else 'tc'   # SGNN_PREC_FP32 / tensor-core scorer
        self.n_params = int(lib().sgnn_param_count(h))
        self.emb_dim = int(lib().sgnn_embed_dim(h))
        self.node_dim = int(lib().sgnn_node_dim(h))
        self.launches = 0  # kernels launched through this object (bench.py's gpu_launches)

    def close(self):
        if getattr(self, 'h', None):
            lib().sgnn_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- parameters ----------------------------------------------------------------
    def param_table(self):
        n = lib().sgnn_param_table(self.h, None, 0)
        arr = (native.ParamInfo * n)()
        lib().sgnn_param_table(self.h, arr, n)
        return [dict(layer=p.layer, name=p.name.decode(), shape=tuple(p.shape[:p.ndim]),
                     offset=int(p.offset)) for p in arr]

    def set_params(self, flat):
        flat = np.ascontiguousarray(flat, np.float32)
        assert flat.size == self.n_params
        check(self.h, lib().sgnn_set_params(self.h, _ptr(flat)))

    def get_params(self):
        out = np.empty(self.n_params, np.float32)
        check(self.h, lib().sgnn_get_params(self.h, _ptr(out)))
        return out

    def init_params(self, seed):
        check(self.h, lib().sgnn_init_params(self.h, int(seed)))

    def device_buffers(self):
        """(params, grads, adam_m, adam_v) as torch views of the handle-owned buffers."""
        import torch
        ptrs = [C.c_void_p() for _ in range(4)]
        check(self.h, lib().sgnn_device_buffers(self.h, *[C.byref(p) for p in ptrs]))
        out = []
        for p in ptrs:
            class _Arr(object):
                pass
            a = _Arr()
            a.__cuda_array_interface__ = {'shape': (self.n_params,), 'typestr': '<f4',
                                          'data': (p.value, False), 'version': 2}
            out.append(torch.as_tensor(a, device=self.device))
        return out

    def reserve(self, max_graphs=0, max_nodes=0, max_nnz=0, max_pairs=0, max_queries=0, max_db=0):
        check(self.h, lib().sgnn_reserve(self.h, max_graphs, max_nodes, max_nnz, max_pairs,
                                         max_queries, max_db))

    # ---- device-resident compute -----------------------------------------------------
    def embed(self, batch, out=None, stream=None):
        import torch
        if out is None:
            out = torch.empty((batch.n_graphs, self.emb_dim), dtype=torch.float32, device=self.device)
        s = batch.struct()
        check(self.h, lib().sgnn_embed_graphs(self.h, C.byref(s), _ptr(out), _stream_ptr(stream)))
        self.launches += 1
        return out

    def node_embeddings(self, batch, stream=None):
        """Last node-layer output of every node, [total_nodes, node_dim] (needs reserve(max_nodes))."""
        import torch
        d = int(lib().sgnn_node_dim(self.h))
        out = torch.empty((batch.total_nodes, d), dtype=torch.float32, device=self.device)
        s = batch.struct()
        check(self.h, lib().sgnn_node_embeddings(self.h, C.byref(s), _ptr(out), _stream_ptr(stream)))
        self.launches += 1
        return out

    def pair_histogram(self, node_emb, node_off, g1, g2, bins=16, stream=None):
        """SimGNN node-node similarity histogram per pair -> (hist [P, bins] float32, counts int32)."""
        import torch
        n = int(g1.numel())
        hist = torch.empty((n, bins), dtype=torch.float32, device=self.device)
        counts = torch.empty((n, bins), dtype=torch.int32, device=self.device)
        n_graphs, total_nodes = int(node_off.numel()) - 1, int(node_emb.shape[0])
        # large graphs on a tensor-core engine: tiles of 128 node rows per graph (sgnn_hist_pairs_tc.cuh); small graphs
        # would leave most of a tile empty and stay on the CUDA-core kernel
        if (self.precision != 'fp32' and self.node_dim <= 16 and bins <= 16 and n_graphs > 0
                and total_nodes >= 48 * n_graphs):
            check(self.h, lib().sgnn_pair_histogram_tc(self.h, _ptr(node_emb), _ptr(node_off), n_graphs, total_nodes,
                                                       _ptr(g1), _ptr(g2), n, int(bins), _ptr(hist), _ptr(counts),
                                                       _stream_ptr(stream)))
            self.launches += 6
            return hist, counts
        check(self.h, lib().sgnn_pair_histogram(self.h, _ptr(node_emb), _ptr(node_off), _ptr(g1), _ptr(g2), n,
                                                int(bins), _ptr(hist), _ptr(counts), _stream_ptr(stream)))
        self.launches += 1
        return hist, counts

    def score_pairs(self, emb, g1, g2, stream=None):
        import torch
        n = int(g1.numel())
        raw = torch.empty(n, dtype=torch.float32, device=self.device)
        sim = torch.empty(n, dtype=torch.float32, device=self.device)
        check(self.h, lib().sgnn_score_pairs(self.h, _ptr(emb), _ptr(g1), _ptr(g2), n, _ptr(raw),
                                             _ptr(sim), _stream_ptr(stream)))
        self.launches += 1
        return raw, sim

    def forward_pairs(self, batch, g1, g2, stream=None):
        if self.has_head:   # one per-pair kernel (the histogram needs the node rows): sgnn_forward_pairs
            import torch
            n = int(g1.numel())
            raw = torch.empty(n, dtype=torch.float32, device=self.device)
            sim = torch.empty(n, dtype=torch.float32, device=self.device)
            s = batch.struct()
            check(self.h, lib().sgnn_forward_pairs(self.h, C.byref(s), _ptr(g1), _ptr(g2), n, None, _ptr(raw),
                                                   _ptr(sim), _stream_ptr(stream)))
            self.launches += 1
            return raw, sim
        emb = self.embed(batch, stream=stream)
        return self.score_pairs(emb, g1, g2, stream=stream)

    def embed_nodes(self, batch, stream=None):
        """(pooled embeddings [G, emb_dim], last node-level rows [total_nodes, node_dim]) in one pass."""
        import torch
        d = int(lib().sgnn_node_dim(self.h))
        emb = torch.empty((batch.n_graphs, self.emb_dim), dtype=torch.float32, device=self.device)
        rows = torch.empty((batch.total_nodes, d), dtype=torch.float32, device=self.device)
        s = batch.struct()
        check(self.h, lib().sgnn_embed_graphs_nodes(self.h, C.byref(s), _ptr(emb), _ptr(rows), _stream_ptr(stream)))
        self.launches += 1
        return emb, rows

    def score_all_pairs_nodes(self, emb_q, rows_q, batch_q, emb_db, rows_db, batch_db, out=None, stream=None):
        """All-pairs scores of a model with a Histogram head (needs both sides' node rows)."""
        import torch
        m, n = emb_q.shape[0], emb_db.shape[0]
        if out is None:
            out = torch.empty((m, n), dtype=torch.float32, device=self.device)
        check(self.h, lib().sgnn_score_all_pairs_nodes(
            self.h, _ptr(emb_q), _ptr(rows_q), _ptr(batch_q.node_off), batch_q.max_nodes, m, _ptr(emb_db),
            _ptr(rows_db), _ptr(batch_db.node_off), batch_db.max_nodes, n, _ptr(out), out.stride(0),
            _stream_ptr(stream)))
        self.launches += 2
        return out

    def hist_all_pairs(self, rows_q, batch_q, rows_db, batch_db, out=None, stages=7, stream=None):
        """Bin counts [m, n, 16] (uint8) of the Histogram layer for every (query, database) pair: one tensor-core
        GEMM with a binning epilogue (sgnn_hist_all_pairs).  Raises ValueError when the graph sizes are outside
        that kernel (use pair_histogram)."""
        import torch
        m, n = batch_q.n_graphs, batch_db.n_graphs
        if out is None:
            out = torch.zeros((m, n, 16), dtype=torch.uint8, device=self.device)
        rc = lib().sgnn_hist_all_pairs(self.h, _ptr(rows_q), _ptr(batch_q.node_off), batch_q.max_nodes, m,
                                       _ptr(rows_db), _ptr(batch_db.node_off), batch_db.max_nodes, n, _ptr(out),
                                       int(stages), _stream_ptr(stream))
        if rc == 4:
            raise ValueError(lib().sgnn_last_error(self.h).decode())
        check(self.h, rc)
        self.launches += {1: 2, 2: 1, 4: 1}.get(int(stages), 4)
        return out

    def score_all_pairs(self, emb_q, emb_db, out=None, stream=None):
        import torch
        m, n = emb_q.shape[0], emb_db.shape[0]
        if out is None:
            out = torch.empty((m, n), dtype=torch.float32, device=self.device)
        check(self.h, lib().sgnn_score_all_pairs(self.h, _ptr(emb_q), m, _ptr(emb_db), n, _ptr(out),
                                                 out.stride(0), _stream_ptr(stream)))
        self.launches += 2 if (self.precision == 'fp32' or self.has_head) else 4   # coefficient matrix, operand prep (query / database side), scorer
        return out

    STAGE_PREP_QUERIES, STAGE_PREP_DATABASE, STAGE_SCORE = 1, 2, 4

    def score_all_pairs_stages(self, emb_q, emb_db, out, stages, stream=None):
        """Selected stages of score_all_pairs (sgnn_score_all_pairs_stages)."""
        m, n = emb_q.shape[0], emb_db.shape[0]
        check(self.h, lib().sgnn_score_all_pairs_stages(self.h, _ptr(emb_q), m, _ptr(emb_db), n, _ptr(out),
                                                        out.stride(0), int(stages), _stream_ptr(stream)))
        if self.precision == 'fp32' or self.has_head:
            self.launches += bin(int(stages) & 5).count('1')
        else:
            self.launches += bin(int(stages)).count('1') + (int(stages) & 1)   # query prep = coefficient matrix + operand
        return out

    def topk(self, sim, k, stream=None):
        import torch
        m, n = sim.shape
        ids = torch.empty((m, k), dtype=torch.int32, device=self.device)
        vals = torch.empty((m, k), dtype=torch.float32, device=self.device)
        check(self.h, lib().sgnn_topk(self.h, _ptr(sim), m, n, sim.stride(0), k, _ptr(ids), _ptr(vals),
                                      _stream_ptr(stream)))
        self.launches += 1
        return ids, vals

    def topk_merge(self, cand_vals, cand_ids, k, stream=None):
        import torch
        m, c = cand_vals.shape
        ids = torch.empty((m, k), dtype=torch.int32, device=self.device)
        vals = torch.empty((m, k), dtype=torch.float32, device=self.device)
        check(self.h, lib().sgnn_topk_merge(self.h, _ptr(cand_vals), _ptr(cand_ids), m, c, k, _ptr(ids),
                                            _ptr(vals), _stream_ptr(stream)))
        self.launches += 1
        return ids, vals

    # ---- multi-GPU candidate exchange over peer memory (sgnn_peer_*) -------------------
    def peer_create(self, world, rank, m_cap, k_cap):
        """Allocate this rank's exchange buffer; returns its 64-byte CUDA IPC handle."""
        buf = (C.c_uint8 * 64)()
        check(self.h, lib().sgnn_peer_create(self.h, int(world), int(rank), int(m_cap), int(k_cap), buf))
        return bytes(buf)

    def peer_connect(self, handles):
        """handles: the world IPC handles in rank order (own included)."""
        blob = b''.join(handles)
        arr = (C.c_uint8 * len(blob)).from_buffer_copy(blob)
        check(self.h, lib().sgnn_peer_connect(self.h, arr))

    def peer_destroy(self):
        if getattr(self, 'h', None):
            lib().sgnn_peer_destroy(self.h)

    def peer_barrier(self, stream=None):
        check(self.h, lib().sgnn_peer_barrier(self.h, _stream_ptr(stream)))

    def topk_exchange(self, vals, ids, id_offset, out_ids=None, out_vals=None, stream=None):
        """Push the local top-k (LOCAL ids + id_offset) to every peer and merge: global (ids, vals)."""
        import torch
        m, k = vals.shape
        out_ids = torch.empty((m, k), dtype=torch.int32, device=self.device) if out_ids is None else out_ids
        out_vals = torch.empty((m, k), dtype=torch.float32, device=self.device) if out_vals is None else out_vals
        check(self.h, lib().sgnn_topk_exchange(self.h, _ptr(vals), _ptr(ids), m, k, int(id_offset), _ptr(out_ids),
                                               _ptr(out_vals), _stream_ptr(stream)))
        self.launches += 2
        return out_ids, out_vals

    def sort_rows(self, sim, stream=None):
        import torch
        m, n = sim.shape
        ids = torch.empty((m, n), dtype=torch.int32, device=self.device)
        check(self.h, lib().sgnn_sort_rows(self.h, _ptr(sim), m, n, sim.stride(0), _ptr(ids),
                                           _stream_ptr(stream)))
        self.launches += 1
        return ids

    def precision_at_ks(self, true_sim, pred_sim, ks, stream=None):
        """Mean over queries of |true_topk_incl(k) & pred_topk(k)| / k for every k in ks
        (src/metrics.py:14-35); true_sim / pred_sim: [m, n] float32 device tensors, higher =
        more similar."""
        import torch
        m, n = pred_sim.shape
        assert tuple(true_sim.shape) == (m, n) and true_sim.is_contiguous() and pred_sim.is_contiguous()
        ks_d = torch.as_tensor(np.asarray(ks, np.int32), device=self.device)
        out = torch.empty(len(ks), dtype=torch.float64, device=self.device)
        kmax = int(max(ks))
        if kmax <= 64 and n >= 4 * kmax and min(ks) >= 1:
            # small k: two top-k launches + a count kernel instead of two full row sorts (sgnn_precision_from_topk)
            _, tv = self.topk(true_sim, kmax, stream=stream)
            pi, _ = self.topk(pred_sim, kmax, stream=stream)
            check(self.h, lib().sgnn_precision_from_topk(self.h, _ptr(true_sim), true_sim.stride(0), _ptr(tv), _ptr(pi), m,
                                                         kmax, _ptr(ks_d), len(ks), _ptr(out), _stream_ptr(stream)))
            self.launches += 1
            return out
        check(self.h, lib().sgnn_precision_at_k(self.h, _ptr(true_sim), _ptr(pred_sim), m, n, _ptr(ks_d),
                                                len(ks), _ptr(out), _stream_ptr(stream)))
        self.launches += 2
        return out

    def kendall_tau(self, true_sim, pred_sim, stream=None):
        """Per-query Kendall tau-b and the exact pair counts [m, 5] = (concordant, discordant,
        ties only in true, ties only in pred, joint ties)."""
        import torch
        m, n = pred_sim.shape
        assert tuple(true_sim.shape) == (m, n) and true_sim.is_contiguous() and pred_sim.is_contiguous()
        counts = torch.empty((m, 5), dtype=torch.int64, device=self.device)
        tau = torch.empty(m, dtype=torch.float64, device=self.device)
        check(self.h, lib().sgnn_kendall_tau(self.h, _ptr(true_sim), _ptr(pred_sim), m, n, _ptr(counts),
                                             _ptr(tau), _stream_ptr(stream)))
        self.launches += 1
        return tau, counts

    def train_step(self, batch, g1, g2, target_sim, global_batch=None, target_sum=None,
                   target_sqsum=None, loss=None, stream=None):
        """Adds the MSE-part gradient of these pairs into the grad buffer; loss[0] += mse part."""
        import torch
        n = int(g1.numel())
        if global_batch is None:
            global_batch = n
        if target_sum is None:
            t = target_sim.double()
            target_sum, target_sqsum = float(t.sum()), float((t * t).sum())
        if loss is None:
            loss = torch.zeros(4, dtype=torch.float32, device=self.device)
        s = batch.struct()
        check(self.h, lib().sgnn_train_step(self.h, C.byref(s), _ptr(g1), _ptr(g2), _ptr(target_sim), n,
                                            int(global_batch), float(target_sum), float(target_sqsum),
                                            _ptr(loss), _stream_ptr(stream)))
        self.launches += 1
        return loss

    def eval_pairs(self, batch, g1, g2, target_sim=None, global_batch=None, stream=None):
        """Forward only on the listed pairs -> (loss[4] or None, raw, sim)."""
        import torch
        n = int(g1.numel())
        raw = torch.empty(n, dtype=torch.float32, device=self.device)
        sim = torch.empty(n, dtype=torch.float32, device=self.device)
        loss, s1, s2 = None, 0.0, 0.0
        if target_sim is not None:
            loss = torch.zeros(4, dtype=torch.float32, device=self.device)
            t = target_sim.double()
            s1, s2 = float(t.sum()), float((t * t).sum())
        s = batch.struct()
        check(self.h, lib().sgnn_eval_pairs(self.h, C.byref(s), _ptr(g1), _ptr(g2), _ptr(target_sim), n,
                                            int(global_batch or n), s1, s2, _ptr(loss), _ptr(raw), _ptr(sim),
                                            _stream_ptr(stream)))
        self.launches += 1
        return loss, raw, sim

    def apply_adam(self, grad_scale=1.0, loss=None, stream=None):
        check(self.h, lib().sgnn_apply_adam(self.h, float(grad_scale), _ptr(loss), _stream_ptr(stream)))
        self.launches += 1

    def apply_adam_allreduce(self, grad_scale=1.0, loss=None, stream=None):
        """Collective: push the gradient blob to every peer, Adam on the rank-ordered sum (peer_create first)."""
        check(self.h, lib().sgnn_apply_adam_allreduce(self.h, float(grad_scale), _ptr(loss), _stream_ptr(stream)))
        self.launches += 2

    def set_embed_tc(self, enable=True):
        """Opt in to the tcgen05 embedding kernel (sgnn_set_embed_tc) for embed_nodes / the head models' ranking paths."""
        check(self.h, lib().sgnn_set_embed_tc(self.h, 1 if enable else 0))

    def adam_device_step(self, enable=True):
        """Adam's step count on the device (sgnn_adam_device_step): train_step + apply_adam become capturable."""
        check(self.h, lib().sgnn_adam_device_step(self.h, 1 if enable else 0))

    def capture_train_steps(self, batch, feeds, loss, grad_scale=1.0):
        """K optimiser steps as ONE CUDA graph: for every (g1, g2, target_sim, global_batch, target_sum, target_sqsum)
        in ``feeds`` (device tensors that stay alive): zero ``loss``, sgnn_train_step, sgnn_apply_adam.  Returns the
        torch.cuda.CUDAGraph; every ``replay()`` runs the K steps (the Adam step count advances on the device)."""
        import torch
        self.adam_device_step(True)
        g = torch.cuda.CUDAGraph()
        side = torch.cuda.Stream(device=self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):
            with torch.cuda.graph(g, stream=side):
                for g1, g2, t, gb, s1, s2 in feeds:
                    loss.zero_()
                    self.train_step(batch, g1, g2, t, gb, s1, s2, loss=loss)
                    self.apply_adam(grad_scale, loss)
        torch.cuda.current_stream(self.device).wait_stream(side)
        return g

    def set_pair_offset(self, offset, stride=1):
        """Data-parallel training: local pair p is pair offset + p * stride of the global batch (keys the dropout masks)."""
        check(self.h, lib().sgnn_set_pair_offset(self.h, int(offset), int(stride)))

    def set_dropout_stream(self, seed, counter=0):
        """Seed / position of the dropout mask stream (one counter value per train/eval call)."""
        check(self.h, lib().sgnn_set_dropout_stream(self.h, int(seed), int(counter)))

    def zero_grads(self, stream=None):
        check(self.h, lib().sgnn_zero_grads(self.h, _stream_ptr(stream)))

    # ---- host-buffer entry points ----------------------------------------------------
    def forward_pairs_host(self, batch, g1, g2, raw=None, sim=None):
        n = len(g1)
        raw = np.empty(n, np.float32) if raw is None else raw
        sim = np.empty(n, np.float32) if sim is None else sim
        s = batch.struct()
        check(self.h, lib().sgnn_forward_pairs_host(self.h, C.byref(s), _ptr(g1), _ptr(g2), n,
                                                    _ptr(raw), _ptr(sim)))
        self.launches += 2
        return raw, sim

    def _rank_launches(self, db, k):
        """Kernels of one ranking step: 2 embeddings, scorer, (top-k), plus the operand preparation of the scorer
        unless the tile embedding kernels wrote the tensor-core operands themselves (graphs of <= 32 nodes, no head)."""
        n = 3 + (1 if k > 0 else 0)
        small = db is None or db.max_nodes <= 32
        if self.precision == 'fp32' or self.has_head:
            n += 1                      # query-side tables
        else:
            n += 2                      # query side: coefficient matrix + operand
            if not small:
                n += 1                  # database-side operand (otherwise written by the embedding kernel)
        return n

    def rank_device(self, queries, db, k, emb_q=None, emb_db=None, sim=None, ids=None, vals=None, stream=None):
        """The whole ranking step on device-resident batches (sgnn_rank_device): the query branch
        overlaps the database branch.  Returns (ids, vals, sim)."""
        import torch
        m, n, E = queries.n_graphs, db.n_graphs, self.emb_dim
        dev = self.device
        emb_q = torch.empty((m, E), dtype=torch.float32, device=dev) if emb_q is None else emb_q
        emb_db = torch.empty((n, E), dtype=torch.float32, device=dev) if emb_db is None else emb_db
        sim = torch.empty((m, n), dtype=torch.float32, device=dev) if sim is None else sim
        if k > 0:
            ids = torch.empty((m, k), dtype=torch.int32, device=dev) if ids is None else ids
            vals = torch.empty((m, k), dtype=torch.float32, device=dev) if vals is None else vals
        sq, sd = queries.struct(), db.struct()
        check(self.h, lib().sgnn_rank_device(self.h, C.byref(sq), C.byref(sd), int(k), _ptr(emb_q), _ptr(emb_db),
                                             _ptr(sim), sim.stride(0), _ptr(ids), _ptr(vals), _stream_ptr(stream)))
        self.launches += self._rank_launches(db, k)
        return ids, vals, sim

    def rank_sharded(self, queries, db_shard, k, id_offset, emb_q=None, emb_db=None, sim=None, ids=None, vals=None,
                     out_ids=None, out_vals=None, stream=None):
        """Collective (peer_create/connect first): rank this rank's database shard, exchange the candidates over
        peer memory, merge.  One CUDA-graph launch per call once the buffers repeat.  Returns global (ids, vals)."""
        import torch
        m, n, E = queries.n_graphs, db_shard.n_graphs, self.emb_dim
        dev = self.device
        f32, i32 = torch.float32, torch.int32
        emb_q = torch.empty((m, E), dtype=f32, device=dev) if emb_q is None else emb_q
        emb_db = torch.empty((n, E), dtype=f32, device=dev) if emb_db is None else emb_db
        sim = torch.empty((m, n), dtype=f32, device=dev) if sim is None else sim
        ids = torch.empty((m, k), dtype=i32, device=dev) if ids is None else ids
        vals = torch.empty((m, k), dtype=f32, device=dev) if vals is None else vals
        out_ids = torch.empty((m, k), dtype=i32, device=dev) if out_ids is None else out_ids
        out_vals = torch.empty((m, k), dtype=f32, device=dev) if out_vals is None else out_vals
        sq, sd = queries.struct(), db_shard.struct()
        check(self.h, lib().sgnn_rank_sharded(self.h, C.byref(sq), C.byref(sd), int(k), int(id_offset), _ptr(emb_q),
                                              _ptr(emb_db), _ptr(sim), sim.stride(0), _ptr(ids), _ptr(vals),
                                              _ptr(out_ids), _ptr(out_vals), _stream_ptr(stream)))
        self.launches += 2 + self._rank_launches(db_shard, k)   # + push, merge
        return out_ids, out_vals

    def rank_host(self, queries, db, k, ids=None, vals=None, sim=None):
        m = queries.n_graphs
        ids = np.empty((m, k), np.int32) if ids is None else ids
        vals = np.empty((m, k), np.float32) if vals is None else vals
        sq, sd = queries.struct(), db.struct()
        check(self.h, lib().sgnn_rank_host(self.h, C.byref(sq), C.byref(sd), k, _ptr(ids), _ptr(vals),
                                           _ptr(sim)))
        self.launches += self._rank_launches(db, k)
        return ids, vals

    def rank_host_compact(self, q_blob, db_blob, k, ids=None, vals=None, sim=None):
        """sgnn_rank_host on compact blobs (PackedBatch.compact())."""
        if ids is None or vals is None:   # the header is only read when the outputs have to be sized here
            m = int(np.frombuffer(bytes(np.asarray(q_blob[4:8])), np.int32)[0])
            ids = np.empty((m, k), np.int32) if ids is None else ids
            vals = np.empty((m, k), np.float32) if vals is None else vals
        check(self.h, lib().sgnn_rank_host_compact(self.h, _ptr(q_blob), _ptr(db_blob), k, _ptr(ids), _ptr(vals),
                                                   _ptr(sim)))
        self.launches += 2 + self._rank_launches(None, k)   # + 2 expansions (compact blobs: graphs of <= 256 nodes)
        return ids, vals

    def train_step_host(self, batch, g1, g2, target_sim, loss=None):
        loss = np.zeros(2, np.float32) if loss is None else loss
        s = batch.struct()
        check(self.h, lib().sgnn_train_step_host(self.h, C.byref(s), _ptr(g1), _ptr(g2), _ptr(target_sim),
                                                 len(g1), _ptr(loss)))
        self.launches += 2
        return loss
