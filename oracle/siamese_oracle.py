"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's Siamese graph-pair model.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import this file.  The product path
(``graphembedding_b200``) must never import it.

Parity status: the host-side pieces restated here (pre-processing, sampling,
similarity kernel, ranking, metrics) are PINNED against the reference's own
modules executed in the build container (``tests/golden/make_golden.py``).
The model math is pinned against the UNMODIFIED reference model code
(``model/Siamese/{layers,models,model_mse}.py``) executed on a torch-backed
stand-in for the TF-1 API (``oracle/tf1_shim.py``); the arithmetic *inside* TF
ops is un-vendored third-party code (TensorFlow ~1.8, ``docker/Dockerfile:1``)
and is restated from its published definition => "parity unpinned at the TF
boundary".

Every function cites the reference file:line it follows (paths relative to
``/root/reference``).  Written with torch on CPU so that fp64 autograd gives
the gradient oracle; ``dtype`` selects fp32 (TF-faithful) or fp64.
"""
import math

import numpy as np
import torch

# ---------------------------------------------------------------------------
# layer-spec grammar  (model/Siamese/layers_factory.py:8-34, 129-135)
# ---------------------------------------------------------------------------
NODE_LAYERS = ('GraphConvolution', 'Dense')
POOL_LAYERS = ('Padding', 'Average', 'Attention')
PAIR_LAYERS = ('NTN', 'Dot')


def _bool(s):
    if s == 'True':
        return True
    if s == 'False':
        return False
    raise RuntimeError('Unknown bool string {}'.format(s))


def parse_layer_specs(spec_strings, input_dim):
    """'Name:k=v,k=v' -> list of dicts (layers_factory.py:11-19)."""
    layers = []
    for i, s in enumerate(spec_strings):
        sp = s.split(':')
        name = sp[0]
        info = {}
        if len(sp) > 1:
            assert len(sp) == 2
            for spec in sp[1].split(','):
                k, v = spec.split('=')
                info[k] = v
        if name == 'GraphConvolution':  # layers_factory.py:37-57
            if not 5 <= len(info) <= 6:
                raise RuntimeError('GraphConvolution layer must have 3-4 specs')
            ind = info.get('input_dim')
            if not ind:
                if i != 0:
                    raise RuntimeError('The input dim for layer {} must be specified'.format(i))
                ind = input_dim
            layers.append(dict(kind=name, input_dim=int(ind), output_dim=int(info['output_dim']),
                               dropout=_bool(info['dropout']), sparse_inputs=_bool(info['sparse_inputs']),
                               act=info['act'], bias=_bool(info['bias'])))
        elif name == 'Average':  # :60-63
            if len(info) != 0:
                raise RuntimeError('Average layer must have 0 specs')
            layers.append(dict(kind=name))
        elif name == 'Attention':  # build extension (layers.py:143-160; F2)
            layers.append(dict(kind=name, input_dim=int(info['input_dim'])))
        elif name == 'NTN':  # :66-74
            if len(info) != 5:
                raise RuntimeError('Average layer must have 0 specs')
            layers.append(dict(kind=name, input_dim=int(info['input_dim']),
                               feature_map_dim=int(info['feature_map_dim']),
                               dropout=_bool(info['dropout']), inneract=info['inneract'],
                               bias=_bool(info['bias'])))
        elif name == 'Dot':  # :77-80
            if len(info) != 0:
                raise RuntimeError('Dot layer must have 0 specs')
            layers.append(dict(kind=name))
        elif name == 'Dense':  # :83-91
            if len(info) != 5:
                raise RuntimeError('Dot layer must have 5 specs')
            layers.append(dict(kind=name, input_dim=int(info['input_dim']),
                               output_dim=int(info['output_dim']), dropout=_bool(info['dropout']),
                               act=info['act'], bias=_bool(info['bias'])))
        elif name == 'Histogram':  # build extension (SimGNN's pairwise node-node histogram)
            if len(info) != 1:
                raise RuntimeError('Histogram layer must have 1 spec')
            layers.append(dict(kind=name, bins=int(info['bins'])))
        elif name == 'Padding':  # :94-98
            if len(info) != 2:
                raise RuntimeError('Padding layer must have 2 specs')
            layers.append(dict(kind=name, max_in_dims=int(info['max_in_dims']),
                               padding_value=int(info['padding_value'])))
        else:
            raise RuntimeError('Unknown layer {}'.format(name))
    return layers


SIMGNN_LAYERS = [  # BASELINE.json configs[0]: SimGNN-style chain (build-chosen widths, SURVEY section 8 notation)
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

DEFAULT_LAYERS = [  # model/Siamese/config.py:44-66
    'GraphConvolution:output_dim=32,act=relu,dropout=True,bias=True,sparse_inputs=True',
    'GraphConvolution:input_dim=32,output_dim=16,act=identity,dropout=True,bias=True,sparse_inputs=False',
    'Dense:input_dim=16,output_dim=1,dropout=True,act=relu,bias=True',
    'Padding:max_in_dims=10,padding_value=0',
    'NTN:input_dim=10,feature_map_dim=10,inneract=relu,dropout=True,bias=True',
]


# ---------------------------------------------------------------------------
# parameters  (layers.py:67-73, 146-151, 172-177, 268-277; inits.py:11-21)
# ---------------------------------------------------------------------------
def param_table(layers):
    """[(layer_idx, var_name, shape)] in the reference's creation order."""
    tab = []
    for li, L in enumerate(layers):
        k = L['kind']
        if k == 'GraphConvolution':
            tab.append((li, 'weights_0', (L['input_dim'], L['output_dim'])))
            if L['bias']:
                tab.append((li, 'bias', (L['output_dim'],)))
        elif k == 'Dense':
            tab.append((li, 'weights', (L['input_dim'], L['output_dim'])))
            if L['bias']:
                tab.append((li, 'bias', (L['output_dim'],)))
        elif k == 'Attention':
            tab.append((li, 'weights', (L['input_dim'], L['input_dim'])))
        elif k == 'NTN':
            D, K = L['input_dim'], L['feature_map_dim']
            tab.append((li, 'weights_W', (D, D, K)))
            tab.append((li, 'weights_V', (K, 2 * D)))
            tab.append((li, 'weights_U', (K, 1)))
            if L['bias']:
                tab.append((li, 'bias', (K,)))
    return tab


def param_count(layers):
    return sum(int(np.prod(s)) for _, _, s in param_table(layers))


def init_params(layers, seed):
    """Glorot-uniform range sqrt(6/(shape[0]+shape[1])) (inits.py:11-15) for
    weights, zeros for biases (inits.py:18-21).  Returns the flat fp32 blob
    (row-major, creation order)."""
    rng = np.random.default_rng(seed)
    parts = []
    for _, name, shape in param_table(layers):
        if name == 'bias':
            parts.append(np.zeros(shape, np.float32).ravel())
        else:
            r = math.sqrt(6.0 / (shape[0] + shape[1]))
            parts.append(rng.uniform(-r, r, size=shape).astype(np.float32).ravel())
    return np.concatenate(parts)


def unflatten(layers, flat, dtype=torch.float64, requires_grad=False):
    """flat blob -> {(layer_idx, name): tensor}."""
    out = {}
    off = 0
    flat = np.asarray(flat)
    for li, name, shape in param_table(layers):
        n = int(np.prod(shape))
        t = torch.tensor(flat[off:off + n].reshape(shape), dtype=dtype)
        if requires_grad:
            t.requires_grad_(True)
        out[(li, name)] = t
        off += n
    assert off == flat.size
    return out


def flatten(layers, pdict):
    return np.concatenate([pdict[(li, name)].detach().cpu().numpy().astype(np.float64).ravel()
                           for li, name, _ in param_table(layers)])


# ---------------------------------------------------------------------------
# host pre-processing  (model/Siamese/graphs.py)
# ---------------------------------------------------------------------------
def normalized_adjacency_coo(n, edges):
    """A_hat = D^-1/2 (A+I) D^-1/2, D = rowsum(A+I), inf -> 0
    (graphs.py:64-76).  Returns (rows, cols, vals float64) in the order the
    reference emits them: the product ``(A D^-1/2)^T D^-1/2`` converted with
    ``.tocoo()`` walks CSR rows of the transposed product, i.e. entries sorted
    by (row, col) of the result (graphs.py:76, 81-85).

    ``edges``: iterable of undirected (u, v) node-index pairs, u != v.
    """
    import scipy.sparse as sp
    a = np.zeros((n, n), np.float64)
    for u, v in edges:
        a[u, v] = 1.0
        a[v, u] = 1.0
    adj = sp.coo_matrix(sp.csr_matrix(a) + sp.eye(n))    # graphs.py:66, 71
    rowsum = np.array(adj.sum(1))                        # :72
    with np.errstate(divide='ignore'):
        dis = np.power(rowsum, -0.5).flatten()           # :73
    dis[np.isinf(dis)] = 0.0                             # :74
    d = sp.diags(dis)                                    # :75
    m = adj.dot(d).transpose().dot(d).tocoo()            # :76  M[i,j] = (A[j,i]*dis[i])*dis[j]
    return m.row.astype(np.int32), m.col.astype(np.int32), m.data.astype(np.float64)


def coo_to_canonical_csr(n, rows, cols, vals):
    """(row, col)-sorted CSR of a COO triple: the order-independent form the
    device packer is compared against bit-for-bit (values as float32, the
    dtype the reference feeds them in, model_mse.py:14-23)."""
    order = np.lexsort((cols, rows))
    r, c, v = np.asarray(rows)[order], np.asarray(cols)[order], np.asarray(vals)[order]
    rowptr = np.zeros(n + 1, np.int32)
    np.add.at(rowptr, r + 1, 1)
    return np.cumsum(rowptr).astype(np.int32), c.astype(np.int32), v.astype(np.float32)


def gaussian_sim(d, yeta):
    """src/similarity.py:55-56."""
    return np.exp(-yeta * np.square(d))


def normalized_dist(d, n1, n2):
    """src/distance.py:59-60."""
    return 2 * d / (n1 + n2)


# ---------------------------------------------------------------------------
# forward  (model/Siamese/layers.py)
# ---------------------------------------------------------------------------
def _act(name, yeta=None):
    if name == 'relu':
        return torch.relu
    if name == 'identity':
        return lambda x: x
    if name == 'sigmoid':
        return torch.sigmoid
    if name == 'tanh':
        return torch.tanh
    if name == 'sim_kernel':  # layers_factory.py:110-112 -> similarity.py:58-60
        return lambda x: torch.exp(-yeta * x * x)
    raise RuntimeError('Unknown activation function {}'.format(name))


# ---------------------------------------------------------------------------
# dropout  (layers.py:45-49, 98-100, 196, 287-288, 332-338)
# ---------------------------------------------------------------------------
# The reference draws its masks from TensorFlow's RNG, which cannot be reproduced without TF.  The
# SEMANTICS are restated here (keep iff floor(keep_prob + u) == 1, kept values scaled by 1/keep_prob;
# the sparse variant on layer 0 masks whole one-hot rows, i.e. one draw per node); the uniform stream
# u is BUILD-DEFINED -- a counter-based hash shared by this oracle and the CUDA kernels, so that masks
# are identical on both sides.  "parity unpinned by reference" for the stream itself.
def _fmix32(x):
    x = np.asarray(x, np.uint64) & np.uint64(0xFFFFFFFF)
    x ^= x >> np.uint64(16)
    x = (x * np.uint64(0x85EBCA6B)) & np.uint64(0xFFFFFFFF)
    x ^= x >> np.uint64(13)
    x = (x * np.uint64(0xC2B2AE35)) & np.uint64(0xFFFFFFFF)
    x ^= x >> np.uint64(16)
    return x


class Dropout(object):
    """Mask generator: ``mask(slot, layer, rows, cols)`` -> float64 array [rows, cols] of
    {0, 1/keep_prob}.  slot = 2 * (pair index) + arm; layer = index in the layer list."""

    def __init__(self, rate, seed=0, counter=0):
        self.rate = float(rate)
        self.keep = np.float32(1.0) - np.float32(rate)
        seed = int(seed)
        s32 = (seed ^ (seed >> 32)) & 0xFFFFFFFF
        self.base = _fmix32(np.uint64(s32) ^ _fmix32((int(counter) + 0x9E3779B9) & 0xFFFFFFFF))

    def uniform(self, slot, layer, rows, cols):
        M = np.uint64(0xFFFFFFFF)
        sk = _fmix32(self.base ^ np.uint64((int(slot) * 0x9E3779B1 + 0x7F4A7C15) & 0xFFFFFFFF))
        lk = _fmix32(sk ^ np.uint64(((int(layer) + 1) * 0x85EBCA77) & 0xFFFFFFFF))
        r = np.arange(rows, dtype=np.uint64)[:, None]
        c = np.arange(cols, dtype=np.uint64)[None, :]
        e = ((r * np.uint64(131) + c) * np.uint64(0x9E3779B9)) & M
        h = _fmix32((lk + e) & M)
        return (h >> np.uint64(8)).astype(np.float32) * np.float32(2.0 ** -24)

    def mask(self, slot, layer, rows, cols):
        if self.rate == 0.0:
            return np.ones((rows, cols))
        u = self.uniform(slot, layer, rows, cols)
        keep = np.floor(self.keep + u) >= 1                                   # layers.py:334-336
        return keep.astype(np.float64) * float(np.float32(1.0) / self.keep)   # :338, tf.nn.dropout


class Graph(object):
    """One pre-processed graph: integer labels (one-hot column per node) and
    the normalised adjacency as COO, as ``ModelGraph`` holds them
    (graphs.py:34-53)."""

    def __init__(self, labels, rows, cols, vals):
        self.labels = np.asarray(labels, np.int64)
        self.n = len(self.labels)
        self.rows = np.asarray(rows, np.int64)
        self.cols = np.asarray(cols, np.int64)
        self.vals = np.asarray(vals, np.float64)

    @staticmethod
    def from_edges(labels, edges):
        r, c, v = normalized_adjacency_coo(len(labels), edges)
        return Graph(labels, r, c, v)

    def a_hat(self, dtype):
        # fed as float32 sparse placeholders (model_mse.py:14-23)
        v = torch.tensor(self.vals.astype(np.float32) if dtype == torch.float32 else self.vals,
                         dtype=dtype)
        a = torch.zeros((self.n, self.n), dtype=dtype)
        a.index_put_((torch.as_tensor(self.rows), torch.as_tensor(self.cols)), v, accumulate=True)
        return a


def node_forward(layers, P, g, input_dim, dtype=torch.float64):
    """Output of the last GraphConvolution / Dense layer, [n, node_dim] (the pooling layer's input)."""
    node_layers = []
    for L in layers:
        if L['kind'] not in NODE_LAYERS:
            break
        node_layers.append(L)
    return graph_forward(node_layers, P, g, input_dim, dtype).reshape(g.n, -1)


def pair_histogram(h1, h2, bins):
    """SimGNN's node-node similarity histogram.  NOT in the reference snapshot (only
    TensorBoard summaries are called histogram there) -- build-defined, shared with the CUDA op:
    s_ij = sigmoid(h1_i . h2_j), fixed-width bins over [0, 1] (tf.histogram_fixed_width clamps the
    top edge into the last bin), normalised by N1 N2.  Returns (hist, counts, scores)."""
    s = torch.sigmoid(h1 @ h2.T).reshape(-1).numpy().astype(np.float64)
    idx = np.minimum(bins - 1, np.floor(s * bins).astype(np.int64))
    counts = np.bincount(idx, minlength=bins).astype(np.int64)
    return counts / float(max(1, s.size)), counts, s


def graph_forward(layers, P, g, input_dim, dtype=torch.float64, drop=None, slot=0, with_nodes=False):
    """Node-level layers + pooling for one graph; returns the flattened
    representation the pair layer consumes (layers.py:91-118, 136-140,
    154-160, 192-205, 223-227).  ``drop``: a Dropout (masks keyed by ``slot``).
    ``with_nodes``: also return the last node-level output [n, node_dim] (the histogram's input)."""
    nodes = None
    def dmask(li, L, x, sparse=False):
        if drop is None or not L.get('dropout', False) or drop.rate == 0.0:
            return x
        if sparse:      # sparse_dropout: one draw per non-zero = per node of the one-hot X
            mk = drop.mask(slot, li, x.shape[0], 1)
        else:
            mk = drop.mask(slot, li, x.shape[0], x.shape[1])
        return x * torch.as_tensor(mk, dtype=dtype)
    a = g.a_hat(dtype)
    x = torch.zeros((g.n, input_dim), dtype=dtype)       # one-hot, row-normalised (graphs.py:55-62)
    x[torch.arange(g.n), torch.as_tensor(g.labels)] = 1.0
    for li, L in enumerate(layers):
        k = L['kind']
        if k == 'GraphConvolution':                       # layers.py:96-118
            x = dmask(li, L, x, sparse=bool(L.get('sparse_inputs', False)))
            pre = x @ P[(li, 'weights_0')]
            out = a @ pre
            if L['bias']:
                out = out + P[(li, 'bias')]
            x = _act(L['act'])(out)
        elif k == 'Dense':                                # :192-205
            x = dmask(li, L, x)
            out = x @ P[(li, 'weights')]
            if L['bias']:
                out = out + P[(li, 'bias')]
            x = _act(L['act'])(out)
        elif k == 'Padding':                              # :223-227
            nodes = x
            if x.shape[0] > L['max_in_dims']:
                raise ValueError('Paddings must be non-negative')
            pad = torch.full((L['max_in_dims'] - x.shape[0], x.shape[1]),
                             float(L['padding_value']), dtype=dtype)
            x = torch.cat([x, pad], 0)
        elif k == 'Average':                              # :136-140
            nodes = x
            x = x.mean(0)
        elif k == 'Attention':                            # :154-160
            nodes = x
            temp = x.mean(0).reshape(1, -1)
            h_avg = torch.tanh((temp @ P[(li, 'weights')]).reshape(-1, 1))
            att = torch.sigmoid(x @ h_avg)
            x = (att.reshape(1, -1) @ x).squeeze()
        elif k in PAIR_LAYERS:
            break
        else:
            raise RuntimeError(k)
    if with_nodes:
        return x.reshape(-1), (x if nodes is None else nodes)
    return x.reshape(-1)


def head_layers(layers):
    """(index of the pair layer, [(index, layer)] of the pair-level layers after it).  A non-empty tail is the
    SimGNN-style head -- BUILD-DEFINED, absent from the reference snapshot: the NTN emits
    its K-vector f_k = U_k * inneract(m_k) (no sum), an optional ``Histogram:bins=B`` appends the pairwise
    node-node histogram of the last node-level outputs (a constant for the gradient, as in SimGNN), and
    ``Dense`` layers (layers.py:192-205 applied to the 1 x (K+B) row) reduce it to the raw score."""
    pi = [i for i, l in enumerate(layers) if l['kind'] in PAIR_LAYERS][0]
    return pi, [(i, l) for i, l in enumerate(layers) if i > pi]


def pair_forward(layers, P, x1, x2, ntn_mode='reference', drop=None, pair=0, h1=None, h2=None):
    """NTN / Dot on one pair of flattened representations (layers.py:249-252,
    282-310).  ``ntn_mode='reference'`` reproduces the (K,1)*(K,1,1)
    broadcast of layers.py:305-308 (F5): score = (sum U)(sum act(m)); 'paper'
    gives sum_k U_k act(m_k).  With a head (see head_layers) h1 / h2 are the node-level outputs."""
    li, L = [(i, l) for i, l in enumerate(layers) if l['kind'] in PAIR_LAYERS][0]
    _, tail = head_layers(layers)
    if L['kind'] == 'Dot':
        return (x1 * x2).sum()
    W, V, U = P[(li, 'weights_W')], P[(li, 'weights_V')], P[(li, 'weights_U')]
    if drop is not None and L.get('dropout', False) and drop.rate != 0.0:    # layers.py:287-288
        x1 = x1 * torch.as_tensor(drop.mask(2 * pair, li, 1, x1.numel()).reshape(-1), dtype=x1.dtype)
        x2 = x2 * torch.as_tensor(drop.mask(2 * pair + 1, li, 1, x2.numel()).reshape(-1), dtype=x2.dtype)
    x1r, x2r = x1.reshape(1, -1), x2.reshape(1, -1)
    fm = []
    for k in range(L['feature_map_dim']):                 # layers.py:295-303
        v_out = V[k].reshape(1, -1) @ torch.cat([x1r.t(), x2r.t()], 0)
        h = ((x1r @ W[:, :, k]) * x2r).sum()
        mid = v_out + h
        if L['bias']:
            mid = mid + P[(li, 'bias')][k]
        fm.append(mid)
    tbp = _act(L['inneract'])(torch.stack(fm))            # (K,1,1)   :305-306
    if tail:
        f = (U.reshape(-1) * tbp.reshape(-1)).reshape(1, -1)
        for hi, HL in tail:
            if HL['kind'] == 'Histogram':
                hist, _, _ = pair_histogram(h1.detach(), h2.detach(), HL['bins'])
                f = torch.cat([f, torch.as_tensor(hist, dtype=f.dtype).reshape(1, -1)], 1)
            else:                                         # Dense on the pair row (layers.py:192-205)
                if drop is not None and HL.get('dropout', False) and drop.rate != 0.0:
                    f = f * torch.as_tensor(drop.mask(2 * pair, hi, 1, f.shape[1]), dtype=f.dtype)
                f = f @ P[(hi, 'weights')]
                if HL['bias']:
                    f = f + P[(hi, 'bias')]
                f = _act(HL['act'])(f)
        assert f.numel() == 1, 'the head must end with Dense:output_dim=1'
        return f.reshape(())
    if ntn_mode == 'reference':
        return (U * tbp).sum()                            # (K,1)*(K,1,1) -> (K,K,1)   :308
    return (U.reshape(-1) * tbp.reshape(-1)).sum()


def forward_pairs(layers, P, g1s, g2s, input_dim, dtype=torch.float64, ntn_mode='reference', drop=None):
    """Raw (pre final_act) scores for B pairs; arm i is paired with arm i+B
    (model_mse.py:117-126)."""
    outs = []
    for p, (g1, g2) in enumerate(zip(g1s, g2s)):
        x1, h1 = graph_forward(layers, P, g1, input_dim, dtype, drop, 2 * p, with_nodes=True)
        x2, h2 = graph_forward(layers, P, g2, input_dim, dtype, drop, 2 * p + 1, with_nodes=True)
        outs.append(pair_forward(layers, P, x1, x2, ntn_mode, drop, p, h1, h2))
    return torch.stack(outs)


def pair_bin_margins(layers, P, g1s, g2s, input_dim, drop=None):
    """Per pair: distance (in score units) of the closest node-node score to a histogram bin edge, or
    inf without a Histogram layer.  A pair whose margin is below fp32 resolution may land in the
    neighbouring bin on the device; tests use this to tell rounding from errors."""
    _, tail = head_layers(layers)
    bins = [L['bins'] for _, L in tail if L['kind'] == 'Histogram']
    out = np.full(len(g1s), np.inf)
    if not bins:
        return out
    for p, (g1, g2) in enumerate(zip(g1s, g2s)):
        _, h1 = graph_forward(layers, P, g1, input_dim, torch.float64, drop, 2 * p, with_nodes=True)
        _, h2 = graph_forward(layers, P, g2, input_dim, torch.float64, drop, 2 * p + 1, with_nodes=True)
        _, _, sc = pair_histogram(h1.detach(), h2.detach(), bins[0])
        out[p] = np.min(np.abs(sc * bins[0] - np.round(sc * bins[0]))) / bins[0]
    return out


def total_loss(layers, P, raw, targets_sim, final_act='sim_kernel', yeta=0.6,
               weight_decay=5e-4, loss_mode='reference'):
    """models.py:67-88 + model_mse.py:145-151.
    reference: (B,1) targets minus (B,) predictions -> (B,B) (F6);
    loss = sum_vars wd*l2_loss(var) + l2_loss(diff)/B, l2_loss(x)=sum(x^2)/2."""
    B = raw.shape[0]
    pred = _act(final_act, yeta)(raw)
    s = torch.as_tensor(np.asarray(targets_sim), dtype=raw.dtype).reshape(B, 1)
    if loss_mode == 'reference':
        diff = s - pred                                   # (B,1)-(B,) -> (B,B)
    else:
        diff = s.reshape(B) - pred
    mse = (diff * diff).sum() / 2 / B
    wd = 0
    for key in P:
        wd = wd + weight_decay * (P[key] * P[key]).sum() / 2
    return wd + mse, mse, pred


class TFAdam(object):
    """tf.train.AdamOptimizer defaults b1=.9 b2=.999 eps=1e-8 (models.py:28-29):
    lr_t = lr*sqrt(1-b2^t)/(1-b1^t); var -= lr_t*m/(sqrt(v)+eps)  (F14)."""

    def __init__(self, n, lr=0.01, b1=0.9, b2=0.999, eps=1e-8):
        self.lr, self.b1, self.b2, self.eps = lr, b1, b2, eps
        self.m = np.zeros(n, np.float64)
        self.v = np.zeros(n, np.float64)
        self.t = 0

    def step(self, params, grads):
        self.t += 1
        lr_t = self.lr * math.sqrt(1 - self.b2 ** self.t) / (1 - self.b1 ** self.t)
        self.m = self.b1 * self.m + (1 - self.b1) * grads
        self.v = self.b2 * self.v + (1 - self.b2) * grads * grads
        return params - lr_t * self.m / (np.sqrt(self.v) + self.eps)


def loss_and_grads(layers, flat_params, g1s, g2s, targets_sim, input_dim, dtype=torch.float64,
                   final_act='sim_kernel', yeta=0.6, weight_decay=5e-4,
                   loss_mode='reference', ntn_mode='reference', drop=None):
    """-> (loss, mse_part, predictions, flat gradient of the TOTAL loss)."""
    P = unflatten(layers, flat_params, dtype, requires_grad=True)
    raw = forward_pairs(layers, P, g1s, g2s, input_dim, dtype, ntn_mode, drop)
    loss, mse, pred = total_loss(layers, P, raw, targets_sim, final_act, yeta, weight_decay, loss_mode)
    keys = [(li, name) for li, name, _ in param_table(layers)]
    grads = torch.autograd.grad(loss, [P[k] for k in keys], allow_unused=True)
    flat_g = np.concatenate([(g if g is not None else torch.zeros_like(P[k])).detach().numpy()
                             .astype(np.float64).ravel() for g, k in zip(grads, keys)])
    return float(loss.detach()), float(mse.detach()), pred.detach().numpy(), flat_g, raw.detach().numpy()


# ---------------------------------------------------------------------------
# ranking + metrics  (src/results.py:51-109, 221-224; src/metrics.py:14-82)
# ---------------------------------------------------------------------------
def sort_id_mat_sim(sim_mat):
    """Descending ranking by a stable ascending mergesort reversed: ties come
    out in DESCENDING index order (results.py:221-224)."""
    return np.argsort(sim_mat, kind='mergesort')[:, ::-1]


def sort_id_mat_dist(dist_mat):
    """results.py:143-144: ascending stable mergesort (ties ascending index)."""
    return np.argsort(dist_mat, kind='mergesort')


def top_k_ids(sort_ids_row, score_row, k, inclusive):
    """results.py:66-82."""
    n = len(sort_ids_row)
    if k < 0 or k >= n:
        raise RuntimeError('Invalid k {}'.format(k))
    if not inclusive:
        return sort_ids_row[:k]
    while k < n:
        if score_row[sort_ids_row[k - 1]] == score_row[sort_ids_row[k]]:
            k += 1
        else:
            break
    return sort_ids_row[:k]


def ranking(sort_ids_row, score_row, gid, one_based=True):
    """results.py:94-109."""
    fid = int(np.where(sort_ids_row == gid)[0][0])
    while fid > 0:
        if score_row[sort_ids_row[fid - 1]] == score_row[sort_ids_row[fid]]:
            fid -= 1
        else:
            break
    return fid + 1 if one_based else fid


def precision_at_ks(true_sort, true_score, pred_sort, pred_score, ks):
    """metrics.py:14-35: truth top-k tie-inclusive, prediction not."""
    m = true_sort.shape[0]
    ps = np.zeros((m, len(ks)))
    for i in range(m):
        for ki, k in enumerate(ks):
            t = top_k_ids(true_sort[i], true_score[i], k, True)
            p = top_k_ids(pred_sort[i], pred_score[i], k, False)
            ps[i][ki] = len(set(t.tolist()).intersection(set(p.tolist()))) / k
    return np.mean(ps, axis=0)


def mean_reciprocal_rank(true_sort, true_score, pred_sort, pred_score):
    """metrics.py:38-64: 1/hmean(min predicted rank over the true top-1 tie set)."""
    m = true_sort.shape[0]
    ranks = np.zeros(m)
    for i in range(m):
        t = top_k_ids(true_sort[i], true_score[i], 1, True)
        ranks[i] = min(ranking(pred_sort[i], pred_score[i], tid) for tid in t)
    return 1.0 / (m / np.sum(1.0 / ranks))


def frobenius_mse(true_sim, pred_sim):
    """metrics.py:67-82: ||S_true - S_pred||_F / (m n)."""
    m, n = true_sim.shape
    return np.linalg.norm(true_sim - pred_sim) / (m * n)


def kendall_tau_counts(x, y):
    """Build-defined (F3; pin = scipy.stats.kendalltau tau-b).  Exact integer
    counts: (concordant, discordant, ties only in x, ties only in y, joint ties)."""
    x = np.asarray(x)
    y = np.asarray(y)
    dx = np.sign(x[:, None] - x[None, :])
    dy = np.sign(y[:, None] - y[None, :])
    iu = np.triu_indices(len(x), 1)
    dx, dy = dx[iu], dy[iu]
    con = int(np.sum(dx * dy > 0))
    dis = int(np.sum(dx * dy < 0))
    tx = int(np.sum((dx == 0) & (dy != 0)))
    ty = int(np.sum((dx != 0) & (dy == 0)))
    txy = int(np.sum((dx == 0) & (dy == 0)))
    return con, dis, tx, ty, txy


def kendall_tau_b(x, y):
    con, dis, tx, ty, txy = kendall_tau_counts(x, y)
    den = math.sqrt(float(con + dis + tx)) * math.sqrt(float(con + dis + ty))
    return (con - dis) / den if den > 0 else float('nan')


# ---------------------------------------------------------------------------
# samplers  (model/Siamese/samplers.py:19-31)
# ---------------------------------------------------------------------------
class RandomSamplerOracle(object):
    """samplers.py:24-31 on a list of ids (mutated in place, F10)."""

    def __init__(self, ids):
        import random
        self._random = random
        self.gs = ids
        self.idx = 0

    def get_pair(self):
        g1 = self.gs[self.idx]
        self.idx += 1
        if self.idx >= len(self.gs):
            self._random.Random(123).shuffle(self.gs)
            self.idx = 0
        g2 = self.gs[self.idx]
        return g1, g2
