"""The multi-GPU data plane (csrc/sgnn_peer.cuh: peer push / merge / gradient push fused into Adam) on real GPUs:
spawns one process per GPU (tests/multi_gpu_worker.py).  Skipped on boxes with a single GPU -- the world_size-2
gloo tests (test_dist_gloo.py) cover the host logic there."""
import os
import subprocess
import sys

import pytest

torch = pytest.importorskip('torch')
pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason='needs >= 2 GPUs')
def test_peer_exchange_and_dp_adam_on_two_gpus():
    n = min(torch.cuda.device_count(), int(os.environ.get('MULTI_GPUS', '2')))
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', str(n),
           '--master-addr', '127.0.0.1', '--master-port', '29541', os.path.join(ROOT, 'tests', 'multi_gpu_worker.py')]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert r.returncode == 0 and 'MULTI_GPU_WORKER_OK' in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
