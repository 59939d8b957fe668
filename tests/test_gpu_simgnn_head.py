"""GPU parity of the SimGNN-style chain (BASELINE.json configs[0]/[1]: 3 GCN + Attention + NTN + pairwise
node-node histogram + FC head) against the CPU oracle.

The histogram layer and the Dense head after the NTN are BUILD-DEFINED (absent from the reference snapshot,
SURVEY a24 / a8 / F1: "parity unpinned by reference"); the pin is oracle/siamese_oracle.py (head_layers,
pair_forward, pair_histogram).  Tolerances: 1e-4 relative for fp32 predictions / loss, 2e-4 of max |g| for
gradients.  Bin indices are integers: a pair whose closest node-node score lies within 1e-5 of a bin edge may
move one count to the neighbouring bin in fp32, so such pairs (reported by O.pair_bin_margins) are excluded
from the value comparison and the gradient tests assert that their seeded inputs contain none."""
import numpy as np
import pytest

from util import CONFIGS, O, corpus, encoder_for, oracle_graphs

torch = pytest.importorskip('torch')
pytestmark = pytest.mark.gpu

RTOL = 1e-4
EDGE = 1e-5


def _setup(cfg_name, num_graphs, n_hi=10, seed=3, **kw):
    from graphembedding_b200.engine import SgnnEngine
    from graphembedding_b200.graphs import pack_nx_graphs
    from graphembedding_b200.layers_factory import parse_layers
    gs = corpus(num_graphs, 2, n_hi, seed)
    enc = encoder_for(gs)
    layers = parse_layers(CONFIGS[cfg_name], enc.input_dim())
    eng = SgnnEngine(layers, enc.input_dim(), **kw)
    flat = O.init_params(O.parse_layer_specs(CONFIGS[cfg_name], enc.input_dim()), seed=11)
    rng = np.random.default_rng(5)
    flat = flat + (rng.standard_normal(flat.shape) * 0.01).astype(np.float32)   # non-zero biases
    eng.set_params(flat)
    return eng, layers, flat, gs, enc, pack_nx_graphs(gs, enc)


def test_param_table_is_the_oracles():
    eng, layers, flat, gs, enc, batch = _setup('simgnn', 4)
    assert eng.n_params == O.param_count(layers) == 10129
    got = [(p['layer'], p['name'], tuple(p['shape'])) for p in eng.param_table()]
    want = [(li, name, tuple(shape)) for li, name, shape in O.param_table(layers)]
    assert [(a, b) for a, b, _ in got] == [(a, b) for a, b, _ in want]
    for (_, _, gs_), (_, _, ws) in zip(got, want):
        assert int(np.prod(gs_)) == int(np.prod(ws))


@pytest.mark.parametrize('cfg,final_act', [('simgnn', 'sigmoid'), ('simgnn', 'sim_kernel'), ('head_nohist', 'sim_kernel'),
                                           ('head_drop', 'sigmoid')])
@pytest.mark.parametrize('n_hi', [10, 30])
def test_forward_pairs_with_head_matches_oracle(cfg, final_act, n_hi):
    eng, layers, flat, gs, enc, batch = _setup(cfg, 64, n_hi=n_hi, final_act=final_act)
    B = 32
    g1 = torch.arange(0, B, dtype=torch.int32, device=eng.device)
    g2 = torch.arange(B, 2 * B, dtype=torch.int32, device=eng.device)
    raw, sim = eng.forward_pairs(batch.to(eng.device), g1, g2)
    P = O.unflatten(layers, flat, torch.float64)
    og = oracle_graphs(gs, enc)
    ref_raw = O.forward_pairs(layers, P, og[:B], og[B:2 * B], enc.input_dim(), torch.float64).numpy()
    ok = O.pair_bin_margins(layers, P, og[:B], og[B:2 * B], enc.input_dim()) > EDGE
    assert ok.sum() >= B - 2
    scale = max(1.0, np.max(np.abs(ref_raw)))
    assert np.max(np.abs(raw.cpu().numpy() - ref_raw)[ok]) <= RTOL * scale
    ref_sim = O._act(final_act, 0.6)(torch.as_tensor(ref_raw)).numpy()
    assert np.max(np.abs(sim.cpu().numpy() - ref_sim)[ok]) <= RTOL
    with pytest.raises(RuntimeError):       # the embedding-only scorer cannot serve a head model
        eng.score_pairs(eng.embed(batch.to(eng.device)), g1, g2)


@pytest.mark.parametrize('cfg', ['simgnn', 'head_nohist', 'head_drop'])
@pytest.mark.parametrize('loss_mode', ['reference', 'paired'])
def test_head_loss_and_gradients_match_fp64_autograd(cfg, loss_mode):
    eng, layers, flat, gs, enc, batch = _setup(cfg, 40, loss_mode=loss_mode, final_act='sigmoid')
    B = 20
    tgt = np.random.default_rng(2).uniform(0.05, 1.0, size=B).astype(np.float32)
    g1 = torch.arange(0, B, dtype=torch.int32, device=eng.device)
    g2 = torch.arange(B, 2 * B, dtype=torch.int32, device=eng.device)
    loss = eng.train_step(batch.to(eng.device), g1, g2, torch.as_tensor(tgt, device=eng.device))
    grads = eng.device_buffers()[1].cpu().numpy().astype(np.float64)
    og = oracle_graphs(gs, enc)
    P = O.unflatten(layers, flat, torch.float64)
    assert np.all(O.pair_bin_margins(layers, P, og[:B], og[B:2 * B], enc.input_dim()) > EDGE)
    l, mse, pred, ref_g, raw = O.loss_and_grads(layers, flat, og[:B], og[B:2 * B], tgt, enc.input_dim(),
                                                final_act='sigmoid', weight_decay=0.0, loss_mode=loss_mode)
    assert abs(float(loss[0]) - mse) <= RTOL * max(abs(mse), 1e-3)
    gscale = np.max(np.abs(ref_g))
    assert gscale > 0
    assert np.max(np.abs(grads - ref_g)) <= 2e-4 * gscale
    # every parameter of the head and of the NTN (U included) receives gradient
    for p in eng.param_table():
        n = int(np.prod(p['shape']))
        if p['layer'] >= [i for i, L in enumerate(layers) if L['kind'] == 'NTN'][0]:
            assert np.any(grads[p['offset']:p['offset'] + n] != 0), p


@pytest.mark.parametrize('cfg', ['simgnn', 'head_drop'])
def test_head_dropout_same_masks_as_oracle(cfg):
    rate = 0.3
    eng, layers, flat, gs, enc, batch = _setup(cfg, 40, loss_mode='paired', final_act='sigmoid', dropout=rate)
    B = 20
    tgt = np.random.default_rng(12).uniform(0.05, 1.0, size=B).astype(np.float32)
    dev = batch.to(eng.device)
    g1 = torch.arange(0, B, dtype=torch.int32, device=eng.device)
    g2 = torch.arange(B, 2 * B, dtype=torch.int32, device=eng.device)
    og = oracle_graphs(gs, enc)
    P = O.unflatten(layers, flat, torch.float64)
    for counter in (0, 7):
        eng.set_dropout_stream(1234567, counter)
        eng.zero_grads()
        loss = eng.train_step(dev, g1, g2, torch.as_tensor(tgt, device=eng.device))
        grads = eng.device_buffers()[1].cpu().numpy().astype(np.float64)
        drop = O.Dropout(rate, seed=1234567, counter=counter)
        ok = O.pair_bin_margins(layers, P, og[:B], og[B:2 * B], enc.input_dim(), drop) > EDGE
        l, mse, pred, ref_g, raw = O.loss_and_grads(layers, flat, og[:B], og[B:2 * B], tgt, enc.input_dim(),
                                                    final_act='sigmoid', weight_decay=0.0, loss_mode='paired',
                                                    drop=drop)
        if ok.all():
            assert abs(float(loss[0]) - mse) <= RTOL * max(abs(mse), 1e-3), (cfg, counter)
            assert np.max(np.abs(grads - ref_g)) <= 2e-4 * np.max(np.abs(ref_g)), (cfg, counter)
        _, raw_d, _ = eng.eval_pairs(dev, g1, g2)          # consumes counter + 1
        drop2 = O.Dropout(rate, seed=1234567, counter=counter + 1)
        ok2 = O.pair_bin_margins(layers, P, og[:B], og[B:2 * B], enc.input_dim(), drop2) > EDGE
        ref_raw = O.forward_pairs(layers, P, og[:B], og[B:2 * B], enc.input_dim(), drop=drop2).numpy()
        assert np.max(np.abs(raw_d.cpu().numpy() - ref_raw)[ok2]) <= RTOL * max(1.0, np.max(np.abs(ref_raw)))
    # forward_pairs (the deterministic entry point) ignores FLAGS.dropout
    raw_nd, _ = eng.forward_pairs(dev, g1, g2)
    ref_nd = O.forward_pairs(layers, P, og[:B], og[B:2 * B], enc.input_dim()).numpy()
    ok3 = O.pair_bin_margins(layers, P, og[:B], og[B:2 * B], enc.input_dim()) > EDGE
    assert np.max(np.abs(raw_nd.cpu().numpy() - ref_nd)[ok3]) <= RTOL * max(1.0, np.max(np.abs(ref_nd)))


@pytest.mark.parametrize('cfg,n_hi', [('simgnn', 10), ('simgnn', 30), ('head_nohist', 10), ('head_drop', 10)])
def test_all_pairs_with_head_equals_pairwise_and_oracle(cfg, n_hi):
    """The m x n loop of test() (train.py:57-61) for a head model: the all-pairs kernel must agree with
    the per-pair kernel and with the oracle on every cell (bin-edge cells excepted)."""
    eng, layers, flat, gs, enc, batch = _setup(cfg, 150, n_hi=n_hi, final_act='sigmoid')
    from graphembedding_b200.graphs import pack_nx_graphs
    m, n = 21, 129
    qb, db = pack_nx_graphs(gs[:m], enc).to(eng.device), pack_nx_graphs(gs[m:], enc).to(eng.device)
    eng.reserve(max_nodes=batch.total_nodes, max_queries=m, max_db=n)
    eq, rq = eng.embed_nodes(qb)
    ed, rd = eng.embed_nodes(db)
    # one-pass node rows == the stand-alone node-level entry point; embeddings == sgnn_embed_graphs
    assert torch.equal(eq, eng.embed(qb)) and torch.equal(ed, eng.embed(db))
    sim = eng.score_all_pairs_nodes(eq, rq, qb, ed, rd, db).cpu().numpy()
    dev = batch.to(eng.device)
    qi = torch.arange(m, dtype=torch.int32, device=eng.device).repeat_interleave(n)
    di = (m + torch.arange(n, dtype=torch.int32, device=eng.device)).repeat(m)
    _, sim_pw = eng.forward_pairs(dev, qi, di)
    sim_pw = sim_pw.cpu().numpy().reshape(m, n)
    P = O.unflatten(layers, flat, torch.float64)
    og = oracle_graphs(gs, enc)
    g1s = [og[i] for i in range(m) for _ in range(n)]
    g2s = [og[m + j] for _ in range(m) for j in range(n)]
    ok = (O.pair_bin_margins(layers, P, g1s, g2s, enc.input_dim()) > EDGE).reshape(m, n)
    assert ok.mean() > 0.99
    ref = torch.sigmoid(O.forward_pairs(layers, P, g1s, g2s, enc.input_dim(), torch.float64)).numpy().reshape(m, n)
    assert np.max(np.abs(sim - ref)[ok]) <= RTOL
    assert np.max(np.abs(sim_pw - ref)[ok]) <= RTOL
    assert np.max(np.abs(sim - sim_pw)[ok]) <= RTOL
    if any(L['kind'] == 'Histogram' for L in layers):
        with pytest.raises(RuntimeError):   # embeddings alone cannot feed the histogram
            eng.score_all_pairs(eq, ed)


def test_rank_entry_points_with_head():
    """sgnn_rank_device / sgnn_rank_host on the SimGNN-style chain: same matrix as the separate calls,
    top-k ids in the reference tie order (results.py:221-224), on the legacy stream and replayed as a graph."""
    eng, layers, flat, gs, enc, batch = _setup('simgnn', 400, final_act='sigmoid', precision='bf16x3')
    from graphembedding_b200.graphs import pack_nx_graphs
    m, n, k = 50, 350, 10
    q, db = pack_nx_graphs(gs[:m], enc), pack_nx_graphs(gs[m:], enc)
    eng.reserve(max_graphs=len(gs), max_nodes=batch.total_nodes, max_nnz=batch.nnz, max_queries=m, max_db=n)
    qd, dbd = q.to(eng.device), db.to(eng.device)
    eq, rq = eng.embed_nodes(qd)
    ed, rd = eng.embed_nodes(dbd)
    sim_ref = eng.score_all_pairs_nodes(eq, rq, qd, ed, rd, dbd)
    ids_ref, vals_ref = eng.topk(sim_ref, k)
    assert np.array_equal(ids_ref.cpu().numpy(), O.sort_id_mat_sim(sim_ref.cpu().numpy())[:, :k])
    ids, vals, sim = eng.rank_device(qd, dbd, k)
    assert torch.equal(sim, sim_ref) and torch.equal(ids, ids_ref) and torch.equal(vals, vals_ref)
    st = torch.cuda.Stream()
    E = eng.emb_dim
    bufs = dict(emb_q=torch.empty((m, E), device=eng.device), emb_db=torch.empty((n, E), device=eng.device),
                sim=torch.empty((m, n), device=eng.device), ids=torch.empty((m, k), dtype=torch.int32, device=eng.device),
                vals=torch.empty((m, k), device=eng.device))
    torch.cuda.synchronize()
    for _ in range(4):      # same buffers on a real stream: eager, capture, replay, replay
        bufs['sim'].zero_()
        torch.cuda.synchronize()
        ids2, vals2, sim2 = eng.rank_device(qd, dbd, k, stream=st, **bufs)
        st.synchronize()
        assert torch.equal(sim2, sim_ref) and torch.equal(ids2, ids_ref) and torch.equal(vals2, vals_ref)
    sim_h = np.empty((m, n), np.float32)
    ids_h, vals_h = eng.rank_host(q.pin(), db.pin(), k, sim=sim_h)
    assert np.array_equal(np.asarray(ids_h), ids_ref.cpu().numpy())
    assert np.array_equal(sim_h, sim_ref.cpu().numpy())


def test_head_chain_validation_errors():
    from graphembedding_b200.engine import SgnnEngine
    from graphembedding_b200.layers_factory import parse_layers
    base = CONFIGS['avg_ntn']
    bad = [
        base + ['Histogram:bins=16'],                                                           # no Dense after it
        base + ['Dense:input_dim=10,output_dim=4,dropout=False,act=relu,bias=True'],            # does not end in 1
        base + ['Dense:input_dim=11,output_dim=1,dropout=False,act=relu,bias=True'],            # width mismatch
        base + ['Histogram:bins=40', 'Dense:input_dim=50,output_dim=1,dropout=False,act=relu,bias=True'],
        base[:3] + ['Dot', 'Dense:input_dim=1,output_dim=1,dropout=False,act=relu,bias=True'],  # head needs an NTN
        base + ['Dense:input_dim=10,output_dim=1,dropout=False,act=relu,bias=True', 'Histogram:bins=4'],
    ]
    for spec in bad:
        with pytest.raises(RuntimeError):
            SgnnEngine(parse_layers(spec, 29), 29)
    with pytest.raises(RuntimeError):
        parse_layers(base + ['Histogram'], 29)          # arity, like layers_factory.py:38-39


@pytest.mark.parametrize('cfg,kw', [('default', dict(precision='bf16x3')), ('default', dict(precision='fp32')),
                                    ('simgnn', dict(final_act='sigmoid'))])
def test_rank_host_compact_equals_rank_host(cfg, kw):
    """The compact wire format (sgnn_compact_batch, expanded on the device) must give the same bits as the
    plain host entry point: the expansion reproduces the CSR arrays exactly."""
    eng, layers, flat, gs, enc, batch = _setup(cfg, 700, **kw)
    from graphembedding_b200.graphs import pack_nx_graphs
    m, n, k = 150, 550, 10
    q, db = pack_nx_graphs(gs[:m], enc), pack_nx_graphs(gs[m:], enc)
    eng.reserve(max_graphs=len(gs), max_nodes=batch.total_nodes, max_nnz=batch.nnz, max_queries=m, max_db=n)
    sim_a, sim_b = np.empty((m, n), np.float32), np.empty((m, n), np.float32)
    ids_a, vals_a = eng.rank_host(q.pin(), db.pin(), k, sim=sim_a)
    qc, dbc = q.compact(), db.compact()
    assert qc is not None and dbc is not None and qc.is_pinned()
    for _ in range(4):      # eager, capture, replay, replay
        sim_b[:] = 0
        ids_b, vals_b = eng.rank_host_compact(qc, dbc, k, sim=sim_b)
        assert np.array_equal(ids_a, ids_b) and np.array_equal(vals_a, vals_b) and np.array_equal(sim_a, sim_b)
