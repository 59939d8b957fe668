"""GPU parity of the tcgen05 embedding kernel (csrc/sgnn_embed_tc.cuh, opt-in through sgnn_set_embed_tc): the dense
GCN layers H W of layers.py:96-118 as 6-product bf16 GEMMs over 128-row tiles of packed graphs, A_hat applied outside
the MMA, pooling (layers.py:136-140, 154-160, 223-227) on the tile.  Pooled embeddings and node rows against the fp64
oracle at the fp32 tolerance, and against the CUDA-core tile kernel at rounding level."""
import numpy as np
import pytest

from util import CONFIGS, O, TOL_FP32, corpus, encoder_for, oracle_graphs

torch = pytest.importorskip('torch')
pytestmark = pytest.mark.gpu

PAD32 = [  # GCN32-GCN16 + Padding: the default-flag pooling on widths the MMA serves
    'GraphConvolution:output_dim=32,act=relu,dropout=False,bias=True,sparse_inputs=True',
    'GraphConvolution:input_dim=32,output_dim=16,act=identity,dropout=False,bias=True,sparse_inputs=False',
    'Padding:max_in_dims=4,padding_value=0',
    'NTN:input_dim=64,feature_map_dim=1,inneract=relu,dropout=False,bias=True',
]


@pytest.mark.parametrize('name,spec,n_hi', [('simgnn', None, 10), ('att3', None, 30), ('avg_ntn', None, 64), ('pad32', PAD32, 4)])
def test_tensor_core_embedding_matches_oracle_and_tile_kernel(name, spec, n_hi):
    from graphembedding_b200.engine import SgnnEngine
    from graphembedding_b200.graphs import pack_nx_graphs
    from graphembedding_b200.layers_factory import parse_layers
    spec = spec or CONFIGS[name]
    kw = dict(final_act='sigmoid') if name == 'simgnn' else {}
    gs = corpus(300, 1, n_hi, 13)
    enc = encoder_for(gs)
    batch = pack_nx_graphs(gs, enc)
    layers = parse_layers(spec, 29)
    flat = O.init_params(O.parse_layer_specs(spec, 29), seed=4)
    flat = flat + (np.random.default_rng(6).standard_normal(flat.shape) * 0.02).astype(np.float32)   # non-zero biases
    out = {}
    for mode in ('tc', 'tile'):
        eng = SgnnEngine(layers, 29, precision='bf16x3', **kw)
        eng.set_params(flat)
        eng.reserve(max_graphs=len(gs), max_nodes=batch.total_nodes, max_nnz=batch.nnz)
        if mode == 'tc':
            eng.set_embed_tc(True)
        emb, rows = eng.embed_nodes(batch.to(eng.device))
        torch.cuda.synchronize()
        out[mode] = (emb.cpu().numpy(), rows.cpu().numpy())
    assert np.max(np.abs(out['tc'][0] - out['tile'][0])) <= 2e-5 * max(1.0, np.max(np.abs(out['tile'][0])))
    assert np.max(np.abs(out['tc'][1] - out['tile'][1])) <= 2e-5 * max(1.0, np.max(np.abs(out['tile'][1])))
    assert not np.array_equal(out['tc'][0], out['tile'][0])          # (the other kernel really ran)
    P = O.unflatten(layers, flat, torch.float64)
    og = oracle_graphs(gs, enc)
    emb = out['tc'][0]
    for i in range(0, len(og), 7):
        ref = O.graph_forward(layers, P, og[i], 29, torch.float64).numpy()
        assert np.max(np.abs(emb[i] - ref)) <= TOL_FP32 * max(1.0, np.max(np.abs(ref))), (name, i)


def test_tensor_core_embedding_is_refused_on_a_cuda_core_engine():
    from graphembedding_b200.engine import SgnnEngine
    from graphembedding_b200.layers_factory import parse_layers
    eng = SgnnEngine(parse_layers(CONFIGS['simgnn'], 29), 29, final_act='sigmoid')
    with pytest.raises(RuntimeError):
        eng.set_embed_tc(True)
