"""GPU parity of the tensor-core histogram of LISTED pairs of large graphs (csrc/sgnn_hist_pairs_tc.cuh, C ABI
sgnn_pair_histogram_tc; Web-shaped graphs, BASELINE configs[4]).

The Histogram layer is BUILD-DEFINED ("parity unpinned by reference"); the pin is the fp64
restatement oracle/siamese_oracle.py: pair_histogram.  Counts are integers: a pair's counts must EQUAL the oracle's
except for the cells that sit within EDGE (in sigmoid units) of a bin edge, each of which may land in the neighbouring
bin (6-product bf16 split accumulated in fp32 + a 4-ulp upward bias: ~5e-7); the test bounds the L1 difference of
every pair by twice its number of such cells and demands exact equality for the pairs that have none."""
import types

import numpy as np
import pytest

from util import CONFIGS, O

torch = pytest.importorskip('torch')
pytestmark = pytest.mark.gpu

EDGE = 2e-6


def _engine(cfg_name, precision):
    from graphembedding_b200.engine import SgnnEngine
    from graphembedding_b200.layers_factory import parse_layers
    eng = SgnnEngine(parse_layers(CONFIGS[cfg_name], 29), 29, precision=precision)
    eng.set_params(O.init_params(O.parse_layer_specs(CONFIGS[cfg_name], 29), seed=3))
    return eng


def _oracle(rows, off, g1, g2, bins):
    want = np.zeros((len(g1), bins), np.int64)
    band = np.zeros(len(g1), np.int64)
    r64 = rows.astype(np.float64)
    for p, (a, b) in enumerate(zip(g1, g2)):
        A, B = r64[off[a]:off[a + 1]], r64[off[b]:off[b + 1]]
        if A.size == 0 or B.size == 0:
            continue
        with np.errstate(over='ignore'):
            s = 1.0 / (1.0 + np.exp(-(A @ B.T)))
        want[p] = np.bincount(np.minimum(bins - 1, np.floor(s * bins).astype(np.int64)).reshape(-1), minlength=bins)
        edge = np.clip(np.round(s * bins), 1, bins - 1)
        band[p] = int((np.abs(s * bins - edge) / bins <= EDGE).sum())
    return want, band


@pytest.mark.parametrize('cfg,scale', [('simgnn', 0.35), ('head_drop', 0.6), ('simgnn', 4.0)])
def test_pair_histogram_tc_counts_match_the_oracle(cfg, scale):
    eng = _engine(cfg, 'bf16x3')
    eng32 = _engine(cfg, 'fp32')
    d = eng.node_dim
    bins = 16 if cfg == 'simgnn' else 7
    rng = np.random.default_rng(11)
    sizes = np.array([1, 50, 127, 128, 129, 300, 500, 64, 257, 0, 384, 100], np.int64)
    off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    rows = (rng.standard_normal((int(off[-1]), d)) * scale).astype(np.float32)
    rows[off[3]:off[3] + 40] = 0.0            # zero rows: every cell with them sits exactly ON the sigmoid(0) edge (goes up)
    if scale > 1:
        rows[off[5]:off[5] + 8] *= 60.0       # |s| far beyond the exponent clamp: sigma == 0 / == 1 cells
    G = len(sizes)
    g1, g2 = np.meshgrid(np.arange(G), np.arange(G), indexing='ij')
    g1, g2 = g1.reshape(-1).astype(np.int32), g2.reshape(-1).astype(np.int32)      # every ordered pair, incl. the empty graph
    dev = eng.device
    rows_d, off_d = torch.as_tensor(rows, device=dev), torch.as_tensor(off, device=dev)
    g1_d, g2_d = torch.as_tensor(g1, device=dev), torch.as_tensor(g2, device=dev)
    assert int(off[-1]) >= 48 * G             # the engine routes this shape to the tensor-core kernel
    launches = eng.launches
    hist, counts = eng.pair_histogram(rows_d, off_d, g1_d, g2_d, bins=bins)
    assert eng.launches - launches == 6       # ... and did
    hist32, counts32 = eng32.pair_histogram(rows_d, off_d, g1_d, g2_d, bins=bins)
    torch.cuda.synchronize()
    got, got32 = counts.cpu().numpy().astype(np.int64), counts32.cpu().numpy().astype(np.int64)
    cells = sizes[g1] * sizes[g2]
    assert np.array_equal(got.sum(1), cells)
    want, band = _oracle(rows, off, g1, g2, bins)
    zero_pair = (g1 == 3) | (g2 == 3)         # exact-edge cells by construction: must go UP, compare with that convention
    diff = np.abs(got - want).sum(1)
    ok = ~zero_pair
    assert np.all(diff[ok] <= 2 * band[ok])
    exact = ok & (band == 0)
    assert exact.sum() >= 20 and np.array_equal(got[exact], want[exact])
    assert np.all(np.abs(got32 - want).sum(1)[ok] <= 2 * band[ok])
    # the normalised histogram is counts / cells
    h = hist.cpu().numpy()
    nz = cells > 0
    assert np.allclose(h[nz], got[nz] / cells[nz, None], rtol=0, atol=1e-7)
    assert np.all(h[~nz] == 0)
    # zero rows: sigmoid(0) = 1/2 goes up, like the CUDA-core kernel
    zz = (g1 == 3) & (g2 == 3)
    assert np.array_equal(got[zz], got32[zz])
    # a second call with other pairs reuses the workspace and overwrites the counts
    hist2, counts2 = eng.pair_histogram(rows_d, off_d, g2_d, g1_d, bins=bins)
    torch.cuda.synchronize()
    assert np.array_equal(counts2.cpu().numpy().reshape(G, G, bins), got.reshape(G, G, bins).transpose(1, 0, 2))


def test_pair_histogram_tc_rejects_what_the_kernel_does_not_serve():
    from graphembedding_b200.native import lib
    eng32 = _engine('simgnn', 'fp32')
    dev = eng32.device
    rows = torch.zeros((256, eng32.node_dim), device=dev)
    off = torch.as_tensor(np.array([0, 128, 256], np.int32), device=dev)
    g = torch.zeros(1, dtype=torch.int32, device=dev)
    out = torch.zeros((1, 16), dtype=torch.int32, device=dev)
    rc = lib().sgnn_pair_histogram_tc(eng32.h, rows.data_ptr(), off.data_ptr(), 2, 256, g.data_ptr(), g.data_ptr(), 1, 16,
                                      None, out.data_ptr(), None)
    assert rc == 4                           # CUDA-core precision: the caller keeps sgnn_pair_histogram
    eng = _engine('simgnn', 'bf16x3')
    rc = lib().sgnn_pair_histogram_tc(eng.h, rows.data_ptr(), off.data_ptr(), 2, 256, g.data_ptr(), g.data_ptr(), 1, 32,
                                      None, out.data_ptr(), None)
    assert rc == 4                           # more bins than nibbles in a register
