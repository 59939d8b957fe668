"""TEST INFRASTRUCTURE ONLY -- batched torch-CPU form of the oracle, used as the timed CPU
baseline (bench.py ``cpu_baseline`` / ``--impl reference``) and cross-checked against the
per-graph restatement (oracle/siamese_oracle.py) in tests/test_oracle_batched.py.

The reference itself is un-batched (one TF sub-graph per input graph, one ``sess.run`` per
test pair, train.py:57-61); this is the "best-effort batched, all host cores" baseline of
BASELINE.md section 4 -- a STRONGER baseline than the reference's own control flow.
Same math as layers.py:91-118, 136-140, 154-160, 192-205, 223-227, 282-310."""
import numpy as np
import torch

from . import siamese_oracle as O


def pad_graphs(graphs, input_dim, dtype=torch.float32):
    """-> labels [G,N] (long, -1 pad), A_hat [G,N,N], mask [G,N], n [G]."""
    G = len(graphs)
    N = max(g.n for g in graphs)
    lab = torch.full((G, N), -1, dtype=torch.long)
    A = torch.zeros((G, N, N), dtype=dtype)
    n = torch.zeros(G, dtype=torch.long)
    for i, g in enumerate(graphs):
        lab[i, :g.n] = torch.as_tensor(g.labels)
        A[i, torch.as_tensor(g.rows), torch.as_tensor(g.cols)] = torch.as_tensor(g.vals, dtype=dtype)
        n[i] = g.n
    mask = (lab >= 0)
    return lab, A, mask, n


def pad_from_packed(batch, dtype=torch.float32):
    """Same, from a graphembedding_b200 PackedBatch's host arrays (no product code runs)."""
    no, rp = np.asarray(batch.node_off), np.asarray(batch.rowptr)
    G, N = batch.n_graphs, int(batch.max_nodes)
    counts = np.diff(no)
    lab = np.full((G, N), -1, np.int64)
    gi = np.repeat(np.arange(G), counts)
    li = np.arange(len(gi)) - np.repeat(no[:-1], counts)
    lab[gi, li] = np.asarray(batch.labels)
    A = np.zeros((G, N, N), np.float32)
    rows = np.repeat(np.arange(len(gi)), np.diff(rp))
    A[gi[rows], li[rows], np.asarray(batch.col)] = np.asarray(batch.val)
    lab_t = torch.as_tensor(lab)
    return lab_t, torch.as_tensor(A).to(dtype), lab_t >= 0, torch.as_tensor(counts.astype(np.int64))


def embed(layers, P, lab, A, mask, n, dtype=torch.float32, with_nodes=False):
    """Pooled representations [G, emb]; with_nodes: also the last node-level rows [G, N, d] (zero beyond n)."""
    mf = mask.to(dtype).unsqueeze(-1)
    x = None
    for li, L in enumerate(layers):
        k = L['kind']
        if k == 'GraphConvolution':
            if x is None:  # one-hot input: row gather
                pre = P[(li, 'weights_0')][lab.clamp(min=0)] * mf
            else:
                pre = x @ P[(li, 'weights_0')]
            out = torch.bmm(A, pre)
            if L['bias']:
                out = out + P[(li, 'bias')]
            x = O._act(L['act'])(out) * mf
        elif k == 'Dense':
            out = x @ P[(li, 'weights')]
            if L['bias']:
                out = out + P[(li, 'bias')]
            x = O._act(L['act'])(out) * mf
        elif k == 'Padding':
            G, N, d = x.shape
            M = L['max_in_dims']
            if int(n.max()) > M:
                raise ValueError('Paddings must be non-negative')
            out = torch.full((G, M, d), float(L['padding_value']), dtype=dtype)
            keep = min(N, M)
            out[:, :keep] = torch.where(mask[:, :keep].unsqueeze(-1), x[:, :keep], out[:, :keep])
            return out.reshape(G, M * d)
        elif k == 'Average':
            out = x.sum(1) / n.to(dtype).unsqueeze(-1)
            return (out, x) if with_nodes else out
        elif k == 'Attention':
            mean = x.sum(1) / n.to(dtype).unsqueeze(-1)
            c = torch.tanh(mean @ P[(li, 'weights')])
            att = torch.sigmoid(torch.einsum('gnd,gd->gn', x, c)) * mask.to(dtype)
            out = torch.einsum('gn,gnd->gd', att, x)
            return (out, x) if with_nodes else out
    raise RuntimeError('no pooling layer')


def _pair_layer(layers):
    return [(i, l) for i, l in enumerate(layers) if l['kind'] in O.PAIR_LAYERS][0]


def score_pairs(layers, P, x1, x2, ntn_mode='reference'):
    li, L = _pair_layer(layers)
    if L['kind'] == 'Dot':
        return (x1 * x2).sum(-1)
    W, V, U = P[(li, 'weights_W')], P[(li, 'weights_V')], P[(li, 'weights_U')].reshape(-1)
    D = x1.shape[1]
    m = torch.einsum('pa,abk,pb->pk', x1, W, x2) + x1 @ V[:, :D].t() + x2 @ V[:, D:].t()
    if L['bias']:
        m = m + P[(li, 'bias')]
    t = O._act(L['inneract'])(m)
    return U.sum() * t.sum(-1) if ntn_mode == 'reference' else t @ U


def score_all_pairs(layers, P, xq, xd, ntn_mode='reference', final_act='sim_kernel', yeta=0.6):
    li, L = _pair_layer(layers)
    if L['kind'] == 'Dot':
        raw = xq @ xd.t()
    else:
        W, V, U = P[(li, 'weights_W')], P[(li, 'weights_V')], P[(li, 'weights_U')].reshape(-1)
        D = xq.shape[1]
        Pq = torch.einsum('qa,abk->qkb', xq, W) + V[:, D:].unsqueeze(0)     # [Q,K,D]
        c = xq @ V[:, :D].t()
        if L['bias']:
            c = c + P[(li, 'bias')]
        m = torch.einsum('qkb,gb->qkg', Pq, xd) + c.unsqueeze(-1)
        t = O._act(L['inneract'])(m)
        raw = U.sum() * t.sum(1) if ntn_mode == 'reference' else torch.einsum('k,qkg->qg', U, t)
    return O._act(final_act, yeta)(raw)


def score_all_pairs_head(layers, P, xq, hq, nq, xd, hd, nd, final_act='sigmoid', yeta=0.6):
    """All-pairs scores of a SimGNN-style chain (oracle head_layers / pair_forward / pair_histogram, batched):
    NTN K-vector f_k = U_k inneract(m_k), node-node histogram of sigmoid(Hq_i . Hd_j) over the real cells,
    Dense head on [f | hist].  hq [Q,Nq,d], hd [M,Nd,d] zero-padded node rows, nq / nd node counts."""
    li, L = _pair_layer(layers)
    _, tail = O.head_layers(layers)
    W, V, U = P[(li, 'weights_W')], P[(li, 'weights_V')], P[(li, 'weights_U')].reshape(-1)
    D = xq.shape[1]
    Pq = torch.einsum('qa,abk->qkb', xq, W) + V[:, D:].unsqueeze(0)
    c = xq @ V[:, :D].t()
    if L['bias']:
        c = c + P[(li, 'bias')]
    m = torch.einsum('qkb,gb->qgk', Pq, xd) + c.unsqueeze(1)
    f = U * O._act(L['inneract'])(m)                                          # [Q,M,K]
    for hi, HL in tail:
        if HL['kind'] == 'Histogram':
            B = HL['bins']
            Q, M = xq.shape[0], xd.shape[0]
            s = torch.sigmoid(torch.einsum('qid,gjd->qgij', hq, hd))
            idx = torch.clamp((s * B).floor().long(), max=B - 1)
            valid = (torch.arange(hq.shape[1]).unsqueeze(0) < nq.unsqueeze(1)).view(Q, 1, -1, 1) & \
                    (torch.arange(hd.shape[1]).unsqueeze(0) < nd.unsqueeze(1)).view(1, M, 1, -1)
            idx = torch.where(valid, idx, torch.full_like(idx, B))               # padding cells -> a discarded bin
            counts = torch.zeros((Q, M, B + 1), dtype=f.dtype)
            counts.scatter_add_(2, idx.reshape(Q, M, -1), torch.ones((Q, M, idx.shape[2] * idx.shape[3]), dtype=f.dtype))
            cells = (nq.view(Q, 1) * nd.view(1, M)).clamp(min=1).to(f.dtype)
            f = torch.cat([f, counts[:, :, :B] / cells.unsqueeze(-1)], -1)
        else:
            f = f @ P[(hi, 'weights')]
            if HL['bias']:
                f = f + P[(hi, 'bias')]
            f = O._act(HL['act'])(f)
    return O._act(final_act, yeta)(f.squeeze(-1))


def train_step(layers, flat, adam, lab, A, mask, n, g1, g2, targets, final_act='sim_kernel', yeta=0.6,
               weight_decay=5e-4, loss_mode='reference', ntn_mode='reference', dtype=torch.float32):
    """One optimizer step (forward, autograd backward, TF-Adam); returns (new flat params, loss)."""
    P = O.unflatten(layers, flat, dtype, requires_grad=True)
    emb = embed(layers, P, lab, A, mask, n, dtype)
    raw = score_pairs(layers, P, emb[g1], emb[g2], ntn_mode)
    loss, _, _ = O.total_loss(layers, P, raw, targets, final_act, yeta, weight_decay, loss_mode)
    keys = [(li, nm) for li, nm, _ in O.param_table(layers)]
    grads = torch.autograd.grad(loss, [P[k] for k in keys])
    g = np.concatenate([t.detach().numpy().astype(np.float64).ravel() for t in grads])
    return adam.step(np.asarray(flat, np.float64), g), float(loss.detach())
