"""ctypes binding of libsgnn.so (include/sgnn.h).  There is no CPU fallback: if the library
is missing it is built with nvcc; if that fails, or no CUDA device is present when a model
is created, the call raises."""
import ctypes as C
import os

import numpy as np

from .csrc import build as _build

SGNN_MAX_LAYERS = 12
ABI_VERSION = 1

LAYER_KINDS = {'GraphConvolution': 0, 'Dense': 1, 'Padding': 2, 'Average': 3, 'Attention': 4,
               'NTN': 5, 'Dot': 6, 'Histogram': 7}
ACTS = {'relu': 0, 'identity': 1, 'sigmoid': 2, 'tanh': 3, 'sim_kernel': 4}
NTN_MODES = {'reference': 0, 'paper': 1}
LOSS_MODES = {'reference': 0, 'paired': 1}
PRECISIONS = {'fp32': 0, 'bf16x3': 1, 'bf16': 2}


class LayerDesc(C.Structure):
    _fields_ = [('kind', C.c_int32), ('input_dim', C.c_int32), ('output_dim', C.c_int32),
                ('act', C.c_int32), ('bias', C.c_int32), ('dropout', C.c_int32),
                ('sparse_inputs', C.c_int32), ('aux', C.c_int32), ('fparam', C.c_float)]


class Config(C.Structure):
    _fields_ = [('abi_version', C.c_int32), ('num_layers', C.c_int32),
                ('layers', LayerDesc * SGNN_MAX_LAYERS), ('input_dim', C.c_int32),
                ('final_act', C.c_int32), ('yeta', C.c_float), ('ntn_mode', C.c_int32),
                ('loss_mode', C.c_int32), ('weight_decay', C.c_float),
                ('learning_rate', C.c_float), ('beta1', C.c_float), ('beta2', C.c_float),
                ('epsilon', C.c_float), ('dropout', C.c_float), ('precision', C.c_int32),
                ('device', C.c_int32)]


class ParamInfo(C.Structure):
    _fields_ = [('layer', C.c_int32), ('name', C.c_char * 16), ('ndim', C.c_int32),
                ('shape', C.c_int32 * 3), ('offset', C.c_int64)]


class GraphBatch(C.Structure):
    _fields_ = [('n_graphs', C.c_int32), ('total_nodes', C.c_int32), ('nnz', C.c_int64),
                ('max_nodes', C.c_int32), ('node_off', C.c_void_p), ('labels', C.c_void_p),
                ('rowptr', C.c_void_p), ('col', C.c_void_p), ('val', C.c_void_p)]


_P = C.c_void_p
_I32, _I64, _F, _D = C.c_int32, C.c_int64, C.c_float, C.c_double
_BP = C.POINTER(GraphBatch)

# name -> (restype, argtypes): every symbol include/sgnn.h declares
SIGNATURES = {
    'sgnn_pack_graphs': (C.c_int, [_I32, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    'sgnn_create': (C.c_int, [C.POINTER(Config), C.POINTER(_P)]),
    'sgnn_destroy': (C.c_int, [_P]),
    'sgnn_last_error': (C.c_char_p, [_P]),
    'sgnn_abi_version': (C.c_int, []),
    'sgnn_param_count': (_I64, [_P]),
    'sgnn_param_table': (C.c_int, [_P, C.POINTER(ParamInfo), _I32]),
    'sgnn_embed_dim': (C.c_int, [_P]),
    'sgnn_set_params': (C.c_int, [_P, _P]),
    'sgnn_get_params': (C.c_int, [_P, _P]),
    'sgnn_init_params': (C.c_int, [_P, C.c_uint64]),
    'sgnn_device_buffers': (C.c_int, [_P, C.POINTER(_P), C.POINTER(_P), C.POINTER(_P), C.POINTER(_P)]),
    'sgnn_reserve': (C.c_int, [_P, _I64, _I64, _I64, _I64, _I64, _I64]),
    'sgnn_embed_graphs': (C.c_int, [_P, _BP, _P, _P]),
    'sgnn_node_dim': (C.c_int, [_P]),
    'sgnn_node_embeddings': (C.c_int, [_P, _BP, _P, _P]),
    'sgnn_embed_graphs_nodes': (C.c_int, [_P, _BP, _P, _P, _P]),
    'sgnn_score_all_pairs_nodes': (C.c_int, [_P, _P, _P, _P, _I32, _I32, _P, _P, _P, _I32, _I32, _P, _I64, _P]),
    'sgnn_hist_all_pairs': (C.c_int, [_P, _P, _P, _I32, _I32, _P, _P, _I32, _I32, _P, _I32, _P]),
    'sgnn_hist_all_pairs_static': (C.c_int, [_I32, _I32, _P, _P, _I32, _I32, _P, _P, _I32, _I32, _P, _P, C.c_char_p, _I32]),
    'sgnn_pair_histogram': (C.c_int, [_P, _P, _P, _P, _P, _I32, _I32, _P, _P, _P]),
    'sgnn_pair_histogram_tc': (C.c_int, [_P, _P, _P, _I32, _I64, _P, _P, _I32, _I32, _P, _P, _P]),
    'sgnn_score_pairs': (C.c_int, [_P, _P, _P, _P, _I32, _P, _P, _P]),
    'sgnn_forward_pairs': (C.c_int, [_P, _BP, _P, _P, _I32, _P, _P, _P, _P]),
    'sgnn_score_all_pairs': (C.c_int, [_P, _P, _I32, _P, _I32, _P, _I64, _P]),
    'sgnn_score_all_pairs_stages': (C.c_int, [_P, _P, _I32, _P, _I32, _P, _I64, _I32, _P]),
    'sgnn_topk': (C.c_int, [_P, _P, _I32, _I32, _I64, _I32, _P, _P, _P]),
    'sgnn_topk_merge': (C.c_int, [_P, _P, _P, _I32, _I32, _I32, _P, _P, _P]),
    'sgnn_sort_rows': (C.c_int, [_P, _P, _I32, _I32, _I64, _P, _P]),
    'sgnn_train_step': (C.c_int, [_P, _BP, _P, _P, _P, _I32, _I32, _D, _D, _P, _P]),
    'sgnn_eval_pairs': (C.c_int, [_P, _BP, _P, _P, _P, _I32, _I32, _D, _D, _P, _P, _P, _P]),
    'sgnn_apply_adam': (C.c_int, [_P, _F, _P, _P]),
    'sgnn_adam_device_step': (C.c_int, [_P, _I32]),
    'sgnn_zero_grads': (C.c_int, [_P, _P]),
    'sgnn_set_dropout_stream': (C.c_int, [_P, C.c_uint64, C.c_uint64]),
    'sgnn_set_pair_offset': (C.c_int, [_P, _I32, _I32]),
    'sgnn_set_embed_tc': (C.c_int, [_P, _I32]),
    'sgnn_precision_at_k': (C.c_int, [_P, _P, _P, _I32, _I32, _P, _I32, _P, _P]),
    'sgnn_precision_from_topk': (C.c_int, [_P, _P, _I64, _P, _P, _I32, _I32, _P, _I32, _P, _P]),
    'sgnn_kendall_tau': (C.c_int, [_P, _P, _P, _I32, _I32, _P, _P, _P]),
    'sgnn_forward_pairs_host': (C.c_int, [_P, _BP, _P, _P, _I32, _P, _P]),
    'sgnn_rank_device': (C.c_int, [_P, _BP, _BP, _I32, _P, _P, _P, _I64, _P, _P, _P]),
    'sgnn_peer_create': (C.c_int, [_P, _I32, _I32, _I32, _I32, _P]),
    'sgnn_peer_connect': (C.c_int, [_P, _P]),
    'sgnn_topk_exchange': (C.c_int, [_P, _P, _P, _I32, _I32, _I32, _P, _P, _P]),
    'sgnn_peer_destroy': (C.c_int, [_P]),
    'sgnn_rank_sharded': (C.c_int, [_P, _BP, _BP, _I32, _I32, _P, _P, _P, _I64, _P, _P, _P, _P, _P]),
    'sgnn_peer_barrier': (C.c_int, [_P, _P]),
    'sgnn_apply_adam_allreduce': (C.c_int, [_P, _F, _P, _P]),
    'sgnn_debug_slice_hints': (C.c_int, [_P, _P, _P]),
    'sgnn_rank_host': (C.c_int, [_P, _BP, _BP, _I32, _P, _P, _P]),
    'sgnn_compact_bound': (_I64, [_BP]),
    'sgnn_compact_batch': (C.c_int, [_BP, _P, _I64, _P]),
    'sgnn_expand_compact_host': (C.c_int, [_P, _P, _P, _P, _P, _P]),
    'sgnn_rank_host_compact': (C.c_int, [_P, _P, _P, _I32, _P, _P, _P]),
    'sgnn_train_step_host': (C.c_int, [_P, _BP, _P, _P, _P, _I32, _P]),
}

_lib = None


def lib():
    """Load (building first if stale) libsgnn.so and bind every exported symbol."""
    global _lib
    if _lib is None:
        path = _build.build()
        l = C.CDLL(path)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(l, name)  # AttributeError if the .so lacks a declared symbol
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


def _ptr(a):
    """numpy array / torch tensor / int / None -> void* value."""
    if a is None:
        return None
    if isinstance(a, int):
        return a
    if isinstance(a, np.ndarray):
        return a.ctypes.data
    return a.data_ptr()  # torch tensor


class SgnnError(RuntimeError):
    pass


def check(handle, rc):
    if rc != 0:
        msg = lib().sgnn_last_error(handle)
        raise SgnnError(msg.decode() if msg else 'libsgnn error {}'.format(rc))


def pack_graphs(node_counts, edge_off, edges):
    """Host packer -> (node_off, rowptr, col, val, max_nodes) numpy arrays."""
    node_counts = np.ascontiguousarray(node_counts, np.int32)
    edge_off = np.ascontiguousarray(edge_off, np.int64)
    edges = np.ascontiguousarray(edges, np.int32).reshape(-1, 2)
    ng = len(node_counts)
    total = int(node_counts.sum())
    cap = total + 2 * len(edges)
    node_off = np.empty(ng + 1, np.int32)
    rowptr = np.empty(total + 1, np.int32)
    col = np.empty(max(cap, 1), np.int32)
    val = np.empty(max(cap, 1), np.float32)
    nnz = C.c_int64(0)
    mx = C.c_int32(0)
    rc = lib().sgnn_pack_graphs(ng, _ptr(node_counts), _ptr(edge_off), _ptr(edges), _ptr(node_off),
                                _ptr(rowptr), _ptr(col), _ptr(val), C.addressof(nnz), C.addressof(mx))
    if rc != 0:
        raise SgnnError('sgnn_pack_graphs failed with code {}'.format(rc))
    return node_off, rowptr, col[:nnz.value].copy(), val[:nnz.value].copy(), int(mx.value)


class PackedBatch(object):
    """A batch of graphs as one block-diagonal CSR, host (numpy) or device (torch) resident."""

    FIELDS = ('node_off', 'labels', 'rowptr', 'col', 'val')

    def __init__(self, node_off, labels, rowptr, col, val, max_nodes):
        self.node_off, self.labels, self.rowptr, self.col, self.val = node_off, labels, rowptr, col, val
        self.max_nodes = int(max_nodes)
        self.n_graphs = int(len(node_off) - 1)
        self.total_nodes = int(len(labels))
        self.nnz = int(len(col))

    def struct(self):
        """The SgnnGraphBatch view of this batch (built once: the arrays of a PackedBatch are never rebound)."""
        st = getattr(self, '_struct', None)
        if st is None:
            st = self._struct = GraphBatch(self.n_graphs, self.total_nodes, self.nnz, self.max_nodes,
                                           _ptr(self.node_off), _ptr(self.labels), _ptr(self.rowptr), _ptr(self.col),
                                           _ptr(self.val))
        return st

    def to(self, device):
        import torch
        return PackedBatch(*[torch.as_tensor(getattr(self, f)).to(device) for f in self.FIELDS],
                           max_nodes=self.max_nodes)

    def pin(self):
        """Pinned host copy whose five arrays lie back to back (16-byte aligned) in ONE allocation, so
        the *_host entry points can upload the batch with a single copy."""
        import torch
        arrs = [np.ascontiguousarray(np.asarray(getattr(self, f))) for f in self.FIELDS]
        offs, total = [], 0
        for a in arrs:
            offs.append(total)
            total += (a.nbytes + 15) // 16 * 16
        arena = torch.empty(max(total, 16), dtype=torch.uint8).pin_memory()
        out = []
        for a, o in zip(arrs, offs):
            t = arena[o:o + a.nbytes].view(torch.float32 if a.dtype == np.float32 else torch.int32)
            t.copy_(torch.as_tensor(a))
            out.append(t)
        pb = PackedBatch(*out, max_nodes=self.max_nodes)
        pb._arena = arena
        return pb

    def compact(self, pinned=True):
        """Compact wire blob (sgnn_compact_batch) as a uint8 torch tensor, pinned by default; None if the batch
        does not fit the byte-wide fields."""
        import torch
        host = PackedBatch(*[np.ascontiguousarray(np.asarray(getattr(self, f))) for f in self.FIELDS],
                           max_nodes=self.max_nodes)
        st = host.struct()
        cap = int(lib().sgnn_compact_bound(C.byref(st)))
        blob = torch.empty(cap, dtype=torch.uint8)
        if pinned and torch.cuda.is_available():
            blob = blob.pin_memory()
        n = C.c_int64(0)
        rc = lib().sgnn_compact_batch(C.byref(st), blob.data_ptr(), cap, C.addressof(n))
        if rc == 4:
            return None
        if rc != 0:
            raise SgnnError('sgnn_compact_batch failed with code {}'.format(rc))
        return blob[:n.value]

    @staticmethod
    def from_compact(blob):
        """Host expansion of a compact blob (sgnn_expand_compact_host) -> PackedBatch (numpy)."""
        hdr = np.frombuffer(bytes(np.asarray(blob[:64])), np.int32)
        G, N, mx = int(hdr[1]), int(hdr[2]), int(hdr[3])
        nnz = int(np.frombuffer(bytes(np.asarray(blob[16:24])), np.int64)[0])
        arr = np.ascontiguousarray(np.asarray(blob))
        node_off, labels = np.empty(G + 1, np.int32), np.empty(N, np.int32)
        rowptr, col, val = np.empty(N + 1, np.int32), np.empty(max(nnz, 1), np.int32), np.empty(max(nnz, 1), np.float32)
        rc = lib().sgnn_expand_compact_host(_ptr(arr), _ptr(node_off), _ptr(labels), _ptr(rowptr), _ptr(col), _ptr(val))
        if rc != 0:
            raise SgnnError('sgnn_expand_compact_host failed with code {}'.format(rc))
        return PackedBatch(node_off, labels, rowptr, col[:nnz], val[:nnz], mx)

    def nbytes(self):
        return 4 * (self.n_graphs + 1 + self.total_nodes + self.total_nodes + 1 + 2 * self.nnz)

    def algorithmic_bytes(self):
        """4N + 4(N+1) + 8 nnz per graph."""
        return 4 * self.total_nodes + 4 * (self.total_nodes + self.n_graphs) + 8 * self.nnz

    @staticmethod
    def from_lists(labels_per_graph, edges_per_graph):
        counts = np.array([len(l) for l in labels_per_graph], np.int32)
        eoff = np.zeros(len(counts) + 1, np.int64)
        for i, e in enumerate(edges_per_graph):
            eoff[i + 1] = eoff[i] + len(e)
        edges = np.concatenate([np.asarray(e, np.int32).reshape(-1, 2) for e in edges_per_graph]) \
            if len(edges_per_graph) and eoff[-1] > 0 else np.zeros((0, 2), np.int32)
        labels = np.concatenate([np.asarray(l, np.int32) for l in labels_per_graph]) \
            if len(counts) else np.zeros(0, np.int32)
        node_off, rowptr, col, val, mx = pack_graphs(counts, eoff, edges)
        return PackedBatch(node_off, labels.astype(np.int32), rowptr, col, val, mx)

    @staticmethod
    def concat(batches):
        node_off = [np.zeros(1, np.int32)]
        rowptr = [np.zeros(1, np.int32)]
        labels, col, val = [], [], []
        no = ro = 0
        for b in batches:
            node_off.append(np.asarray(b.node_off[1:]) + no)
            rowptr.append(np.asarray(b.rowptr[1:]) + ro)
            labels.append(b.labels); col.append(b.col); val.append(b.val)
            no += b.total_nodes
            ro += b.nnz
        return PackedBatch(np.concatenate(node_off).astype(np.int32), np.concatenate(labels).astype(np.int32),
                           np.concatenate(rowptr).astype(np.int32), np.concatenate(col).astype(np.int32),
                           np.concatenate(val).astype(np.float32), max(b.max_nodes for b in batches))

    def select(self, idx):
        """Sub-batch of the listed graphs (host arrays)."""
        idx = np.asarray(idx, np.int64)
        no, rp = np.asarray(self.node_off), np.asarray(self.rowptr)
        counts = (no[idx + 1] - no[idx]).astype(np.int64)
        node_off = np.zeros(len(idx) + 1, np.int32)
        node_off[1:] = np.cumsum(counts)
        node_ids = np.concatenate([np.arange(no[g], no[g + 1]) for g in idx]) if len(idx) else np.zeros(0, np.int64)
        row_nnz = (rp[node_ids + 1] - rp[node_ids]).astype(np.int64)
        rowptr = np.zeros(len(node_ids) + 1, np.int32)
        rowptr[1:] = np.cumsum(row_nnz)
        ent = np.concatenate([np.arange(rp[no[g]], rp[no[g + 1]]) for g in idx]) if len(idx) else np.zeros(0, np.int64)
        mx = int(counts.max()) if len(idx) else 0
        return PackedBatch(node_off, np.asarray(self.labels)[node_ids].astype(np.int32), rowptr,
                           np.asarray(self.col)[ent].astype(np.int32),
                           np.asarray(self.val)[ent].astype(np.float32), mx)


def make_config(layers, input_dim, final_act='sim_kernel', yeta=0.6, ntn_mode='reference',
                loss_mode='reference', weight_decay=5e-4, learning_rate=0.01, dropout=0.0,
                precision='fp32', device=0):
    """``layers``: list of dicts as produced by layers_factory.parse_layers."""
    cfg = Config()
    cfg.abi_version = ABI_VERSION
    if len(layers) > SGNN_MAX_LAYERS:
        raise SgnnError('at most {} layers'.format(SGNN_MAX_LAYERS))
    cfg.num_layers = len(layers)
    for i, L in enumerate(layers):
        d = cfg.layers[i]
        d.kind = LAYER_KINDS[L['kind']]
        d.input_dim = int(L.get('input_dim', 0) or 0)
        d.output_dim = int(L.get('output_dim', 0) or 0)
        d.act = ACTS[L.get('act', L.get('inneract', 'identity'))]
        d.bias = int(bool(L.get('bias', False)))
        d.dropout = int(bool(L.get('dropout', False)))
        d.sparse_inputs = int(bool(L.get('sparse_inputs', False)))
        d.aux = int(L.get('max_in_dims', L.get('feature_map_dim', L.get('bins', 0))) or 0)
        d.fparam = float(L.get('padding_value', 0.0))
    cfg.input_dim = int(input_dim)
    cfg.final_act = ACTS[final_act]
    cfg.yeta = float(yeta)
    cfg.ntn_mode = NTN_MODES[ntn_mode]
    cfg.loss_mode = LOSS_MODES[loss_mode]
    cfg.weight_decay = float(weight_decay)
    cfg.learning_rate = float(learning_rate)
    cfg.beta1, cfg.beta2, cfg.epsilon = 0.9, 0.999, 1e-8
    cfg.dropout = float(dropout)
    cfg.precision = PRECISIONS[precision]
    cfg.device = int(device)
    return cfg
