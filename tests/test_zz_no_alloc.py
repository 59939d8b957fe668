"""The allocation-free histogram entry point (sgnn_hist_all_pairs_static: scratch in a file-scope device array,
no cudaMalloc, no handle) against the existing entry point (sgnn_hist_all_pairs, sgnn_reserve workspace).  No
counterpart in the reference.  Both run the same kernels, so the bin
counts must be EQUAL byte for byte -- on two calls in a row of different sizes (the shared array keeps the first
call's contents).  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from test_gpu_hist_tc import _engine, _side

torch = pytest.importorskip('torch')
pytestmark = pytest.mark.gpu


def _static(d, bins, rq, bq, rd, bd, out):
    from graphembedding_b200.native import lib, _ptr
    err = C.create_string_buffer(512)
    rc = lib().sgnn_hist_all_pairs_static(d, bins, _ptr(rq), _ptr(bq.node_off), bq.max_nodes, bq.n_graphs, _ptr(rd),
                                          _ptr(bd.node_off), bd.max_nodes, bd.n_graphs, _ptr(out), None, err, 512)
    return rc, err.value.decode()


def test_static_scratch_histogram_equals_the_workspace_entry_point_twice():
    eng, layers, flat = _engine('simgnn', precision='bf16x3')
    d, bins = eng.node_dim, 16
    rng = np.random.default_rng(77)
    for call, (m, n, qmax, dbmax) in enumerate([(41, 311, 10, 10), (17, 90, 15, 17)]):   # second call: smaller, other tile shapes
        sq = rng.integers(1, qmax + 1, size=m); sq[0] = qmax
        sd = rng.integers(1, dbmax + 1, size=n); sd[0] = dbmax
        bq, rq, _ = _side(sq, d, rng, 0.8, eng.device)
        bd, rd, od = _side(sd, d, rng, 0.8, eng.device)
        rq_d, rd_d = torch.as_tensor(rq, device=eng.device), torch.as_tensor(rd, device=eng.device)
        eng.reserve(max_nodes=int(od[-1]), max_queries=m, max_db=n)
        want = eng.hist_all_pairs(rq_d, bq, rd_d, bd)
        got = torch.zeros((m, n, 16), dtype=torch.uint8, device=eng.device)
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        rc, msg = _static(d, bins, rq_d, bq, rd_d, bd, got)
        torch.cuda.synchronize()
        assert rc == 0, msg
        if call == 1:   # (the first call may still load kernels lazily, which the driver backs with device memory)
            assert torch.cuda.mem_get_info()[0] == free_before      # nothing was allocated on the device
        assert torch.equal(got, want)
        assert np.array_equal(got.cpu().numpy().astype(np.int64).sum(-1), np.outer(sq, sd))


def test_static_scratch_histogram_rejects_what_does_not_fit():
    eng, layers, flat = _engine('simgnn', precision='bf16x3')
    rng = np.random.default_rng(3)
    bq, rq, _ = _side([5, 4], 16, rng, 1.0, eng.device)
    bd, rd, _ = _side([6] * 8, 16, rng, 1.0, eng.device)
    rq_d, rd_d = torch.as_tensor(rq, device=eng.device), torch.as_tensor(rd, device=eng.device)
    out = torch.zeros((2, 8, 16), dtype=torch.uint8, device=eng.device)
    bd.n_graphs, keep = 4_000_000, bd.n_graphs                       # claims a database far beyond the static array
    rc, msg = _static(16, 16, rq_d, bq, rd_d, bd, out)
    assert rc == 5 and 'SGNN_HIST_STATIC_SCRATCH_BYTES' in msg
    bd.n_graphs = keep
    bq.max_nodes = 16                                                # outside the kernel
    rc, msg = _static(16, 16, rq_d, bq, rd_d, bd, out)
    assert rc == 4 and msg
