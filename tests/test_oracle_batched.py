"""The batched CPU baseline (oracle/batched.py) must agree with the per-graph restatement."""
import numpy as np
import pytest

from util import CONFIGS, O, corpus, encoder_for, oracle_graphs
from oracle import batched as OB
from graphembedding_b200.graphs import pack_nx_graphs

torch = pytest.importorskip('torch')


@pytest.mark.parametrize('cfg', ['default', 'avg_ntn', 'att3', 'dense_dot'])
def test_batched_equals_per_graph(cfg):
    gs = corpus(30, 2, 10, seed=9)
    enc = encoder_for(gs)
    layers = O.parse_layer_specs(CONFIGS[cfg], enc.input_dim())
    flat = O.init_params(layers, 3)
    P = O.unflatten(layers, flat, torch.float64)
    og = oracle_graphs(gs, enc)
    lab, A, mask, n = OB.pad_graphs(og, enc.input_dim(), torch.float64)
    lab2, A2, mask2, n2 = OB.pad_from_packed(pack_nx_graphs(gs, enc), torch.float64)
    assert torch.equal(lab, lab2) and torch.equal(n, n2)
    assert torch.allclose(A, A2, rtol=0, atol=1e-7)          # packed values are float32
    emb = OB.embed(layers, P, lab, A, mask, n, torch.float64)
    for i, g in enumerate(og):
        ref = O.graph_forward(layers, P, g, enc.input_dim())
        assert torch.allclose(emb[i], ref, rtol=1e-10, atol=1e-12)
    raw = OB.score_pairs(layers, P, emb[:15], emb[15:])
    ref = O.forward_pairs(layers, P, og[:15], og[15:], enc.input_dim())
    assert torch.allclose(raw, ref, rtol=1e-10, atol=1e-12)
    sim = OB.score_all_pairs(layers, P, emb[:7], emb[7:])
    for i in range(7):
        for j in (0, 5, 22):
            r = O.pair_forward(layers, P, emb[i], emb[7 + j])
            assert abs(float(sim[i, j]) - float(torch.exp(-0.6 * r * r))) < 1e-10


def test_batched_train_step_equals_oracle():
    gs = corpus(20, 2, 10, seed=10)
    enc = encoder_for(gs)
    layers = O.parse_layer_specs(CONFIGS['default'], enc.input_dim())
    flat = O.init_params(layers, 4).astype(np.float64)
    og = oracle_graphs(gs, enc)
    lab, A, mask, n = OB.pad_graphs(og, enc.input_dim(), torch.float64)
    tgt = np.linspace(0.1, 0.9, 10)
    a1, a2 = O.TFAdam(flat.size), O.TFAdam(flat.size)
    g1, g2 = torch.arange(0, 10), torch.arange(10, 20)
    new, loss = OB.train_step(layers, flat, a1, lab, A, mask, n, g1, g2, tgt, dtype=torch.float64)
    l, _, _, g, _ = O.loss_and_grads(layers, flat, og[:10], og[10:], tgt, enc.input_dim())
    assert abs(l - loss) < 1e-12 and np.allclose(new, a2.step(flat, g), rtol=0, atol=1e-12)


def test_pair_histogram_known_answers():
    """Build-defined SimGNN histogram (a24): zero embeddings -> every score is sigmoid(0) = 0.5, large
    positive dot products land in the LAST bin (top edge clamped), counts sum to N1 N2."""
    import torch
    from util import O
    h1, h2 = torch.zeros((3, 4), dtype=torch.float64), torch.zeros((5, 4), dtype=torch.float64)
    hist, counts, s = O.pair_histogram(h1, h2, 16)
    assert counts[8] == 15 and counts.sum() == 15 and abs(hist.sum() - 1.0) < 1e-12
    h1 = torch.full((2, 4), 10.0, dtype=torch.float64)
    hist, counts, s = O.pair_histogram(h1, h1, 8)
    assert counts[7] == 4 and (s <= 1.0).all()
    hist, counts, s = O.pair_histogram(h1, -h1, 8)
    assert counts[0] == 4


@pytest.mark.parametrize('cfg', ['simgnn', 'head_drop', 'head_nohist'])
def test_batched_head_scores_equal_per_pair_oracle(cfg):
    """The batched SimGNN-style all-pairs port (bench.py's cpu_baseline for the headline model) against the
    per-pair restatement (oracle pair_forward with head + pair_histogram), fp64, every cell."""
    gs = corpus(26, 2, 10, seed=4)
    enc = encoder_for(gs)
    layers = O.parse_layer_specs(CONFIGS[cfg], enc.input_dim())
    flat = O.init_params(layers, 5)
    P = O.unflatten(layers, flat, torch.float64)
    og = oracle_graphs(gs, enc)
    lab, A, mask, n = OB.pad_graphs(og, enc.input_dim(), torch.float64)
    emb, rows = OB.embed(layers, P, lab, A, mask, n, torch.float64, with_nodes=True)
    m = 6
    sim = OB.score_all_pairs_head(layers, P, emb[:m], rows[:m], n[:m], emb[m:], rows[m:], n[m:], final_act='sigmoid')
    g1s = [og[i] for i in range(m) for _ in range(len(og) - m)]
    g2s = [og[m + j] for _ in range(m) for j in range(len(og) - m)]
    ref = torch.sigmoid(O.forward_pairs(layers, P, g1s, g2s, enc.input_dim(), torch.float64)).reshape(m, -1)
    assert torch.allclose(sim, ref, rtol=1e-10, atol=1e-12)
