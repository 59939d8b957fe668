"""CPU tests of the SimGNN-style head (Histogram + Dense layers after the NTN): oracle restatement, layer
grammar and the POD config handed to sgnn_create.  The head is BUILD-DEFINED (absent from the reference
snapshot) -- "parity unpinned by reference"; these tests pin the oracle to known
answers and to finite differences so that the GPU parity tests have a checked target."""
import numpy as np
import pytest

from util import CONFIGS, O, corpus, encoder_for, oracle_graphs

torch = pytest.importorskip('torch')


def _inputs(cfg, num=12, seed=3):
    gs = corpus(num, 2, 10, seed)
    enc = encoder_for(gs)
    layers = O.parse_layer_specs(CONFIGS[cfg], enc.input_dim())
    flat = O.init_params(layers, seed=11) + (np.random.default_rng(5).standard_normal(O.param_count(layers)) * 0.01
                                             ).astype(np.float32)
    return layers, flat, oracle_graphs(gs, enc), enc


def test_simgnn_chain_shape_and_param_count():
    layers, flat, og, enc = _inputs('simgnn')
    assert O.param_count(layers) == 10129          # "SimGNN-style ~10 129"
    pi, tail = O.head_layers(layers)
    assert layers[pi]['kind'] == 'NTN' and [L['kind'] for _, L in tail] == ['Histogram'] + ['Dense'] * 4
    names = [(li, n) for li, n, _ in O.param_table(layers)]
    assert names[-8:] == [(6, 'weights'), (6, 'bias'), (7, 'weights'), (7, 'bias'), (8, 'weights'), (8, 'bias'),
                          (9, 'weights'), (9, 'bias')]


def test_product_grammar_equals_oracle_grammar_and_config():
    from graphembedding_b200.layers_factory import parse_layers
    from graphembedding_b200 import native
    for cfg in ('simgnn', 'head_nohist', 'head_drop'):
        a = parse_layers(CONFIGS[cfg], 29)
        b = O.parse_layer_specs(CONFIGS[cfg], 29)
        assert len(a) == len(b)
        for x, y in zip(a, b):
            assert x == y, (x, y)
    c = native.make_config(parse_layers(CONFIGS['simgnn'], 29), 29, final_act='sigmoid')
    assert c.num_layers == 10
    assert c.layers[5].kind == native.LAYER_KINDS['Histogram'] == 7 and c.layers[5].aux == 16
    assert c.layers[6].kind == native.LAYER_KINDS['Dense'] and (c.layers[6].input_dim, c.layers[6].output_dim) == (32, 16)
    assert c.layers[4].aux == 16 and c.layers[9].output_dim == 1
    with pytest.raises(RuntimeError):
        parse_layers(CONFIGS['avg_ntn'] + ['Histogram:bins=4,foo=1'], 29)


def test_head_forward_known_answer():
    """Hand-computed: NTN K-vector f_k = U_k relu(m_k), 2-bin histogram, one Dense row."""
    spec = ['GraphConvolution:output_dim=2,act=identity,dropout=False,bias=False,sparse_inputs=True',
            'Average',
            'NTN:input_dim=2,feature_map_dim=2,inneract=relu,dropout=False,bias=True',
            'Histogram:bins=2',
            'Dense:input_dim=4,output_dim=1,dropout=False,act=identity,bias=True']
    layers = O.parse_layer_specs(spec, 2)
    W0 = np.array([[1.0, 0.0], [0.0, -2.0]])
    Wn = np.zeros((2, 2, 2)); Wn[0, 0, 0] = 1.0; Wn[1, 1, 1] = 0.5
    V = np.array([[0.5, 0, 0, 0], [0, 0, 0, 1.0]])
    U = np.array([[2.0], [-1.0]]); b = np.array([0.1, -0.2])
    Wd = np.array([[1.0], [2.0], [3.0], [4.0]]); bd = np.array([0.25])
    flat = np.concatenate([a.ravel() for a in (W0, Wn, V, U, b, Wd, bd)])
    g1 = O.Graph.from_edges([0, 1], [(0, 1)])      # A_hat = [[.5,.5],[.5,.5]]
    g2 = O.Graph.from_edges([1], [])               # single node, A_hat = [[1]]
    P = O.unflatten(layers, flat, torch.float64)
    raw = O.forward_pairs(layers, P, [g1], [g2], 2, torch.float64).numpy()[0]
    h1 = np.array([[0.5, -1.0], [0.5, -1.0]]); h2 = np.array([[0.0, -2.0]])
    x1, x2 = h1.mean(0), h2.mean(0)
    m = np.array([x1 @ Wn[:, :, k] @ x2 + V[k] @ np.concatenate([x1, x2]) + b[k] for k in range(2)])
    f = U.ravel() * np.maximum(m, 0)
    s = 1 / (1 + np.exp(-(h1 @ h2.T).ravel()))     # both cells: sigmoid(2) = 0.88 -> bin 1
    hist = np.bincount(np.minimum(1, np.floor(s * 2).astype(int)), minlength=2) / 2.0
    want = np.concatenate([f, hist]) @ Wd.ravel() + bd[0]
    assert hist.tolist() == [0.0, 1.0]
    assert abs(raw - want) < 1e-12


@pytest.mark.parametrize('cfg', ['simgnn', 'head_drop'])
def test_head_gradients_match_finite_differences(cfg):
    layers, flat, og, enc = _inputs(cfg)
    B = 6
    tgt = np.linspace(0.2, 0.9, B)
    drop = O.Dropout(0.3, seed=9, counter=2) if cfg == 'head_drop' else None
    kw = dict(final_act='sigmoid', weight_decay=1e-3, loss_mode='paired', drop=drop)
    l0, _, _, g, _ = O.loss_and_grads(layers, flat.astype(np.float64), og[:B], og[B:2 * B], tgt, 29, **kw)
    rng = np.random.default_rng(0)
    idx = rng.choice(flat.size, 40, replace=False)
    eps = 1e-6
    for i in idx:
        p = flat.astype(np.float64).copy(); p[i] += eps
        lp = O.loss_and_grads(layers, p, og[:B], og[B:2 * B], tgt, 29, **kw)[0]
        p[i] -= 2 * eps
        lm = O.loss_and_grads(layers, p, og[:B], og[B:2 * B], tgt, 29, **kw)[0]
        fd = (lp - lm) / (2 * eps)
        # the histogram is piecewise constant: finite differences see it as a constant too unless a score
        # crosses a bin edge inside +-eps (margins of these inputs are > 1e-4)
        assert abs(fd - g[i]) <= 1e-5 * max(1.0, abs(g[i])) + 1e-8, (i, fd, g[i])


def test_histogram_is_a_constant_for_the_gradient_and_sums_to_one():
    layers, flat, og, enc = _inputs('simgnn')
    P = O.unflatten(layers, flat, torch.float64)
    for a, b in zip(og[:6], og[6:12]):
        _, h1 = O.graph_forward(layers, P, a, 29, torch.float64, with_nodes=True)
        _, h2 = O.graph_forward(layers, P, b, 29, torch.float64, with_nodes=True)
        hist, counts, sc = O.pair_histogram(h1, h2, 16)
        assert counts.sum() == a.n * b.n and abs(hist.sum() - 1.0) < 1e-12
        assert h1.shape == (a.n, 16) and h2.shape == (b.n, 16)
    assert np.all(np.isfinite(O.pair_bin_margins(layers, P, og[:6], og[6:12], 29)))
    assert np.all(np.isinf(O.pair_bin_margins(O.parse_layer_specs(CONFIGS['head_nohist'], 29), P, og[:2], og[2:4], 29)))
