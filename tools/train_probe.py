"""A few DP-training steps of the bench workload (for ncu captures of train_pairs_kernel)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from graphembedding_b200 import distributed as DS
from graphembedding_b200.engine import SgnnEngine
from graphembedding_b200.layers_factory import parse_layers
enc, batch, batches = bench.make_train_data(0, 1)
simgnn = os.environ.get('PROBE_LAYERS') == 'simgnn'
eng = SgnnEngine(parse_layers(bench.SIMGNN_LAYERS if simgnn else bench.TRAIN_LAYERS, 29), 29, loss_mode='paired',
                 ntn_mode='reference', final_act='sigmoid' if simgnn else 'sim_kernel')
eng.init_params(123)
dev = eng.device
bd = batch.to(dev)
feeds = [(torch.as_tensor(a, device=dev), torch.as_tensor(b, device=dev), torch.as_tensor(t, device=dev), s1, s2)
         for a, b, t, s1, s2 in batches]
loss = torch.zeros(4, dtype=torch.float32, device=dev)
grads = eng.device_buffers()[1]
n = int(os.environ.get('PROBE_STEPS', '5'))
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
for i in range(n):
    if i == n // 2: ev[0].record()
    g1, g2, t, s1, s2 = feeds[i % len(feeds)]
    loss.zero_()
    DS.data_parallel_step(eng, bd, g1, g2, t, bench.TRAIN_BATCH, s1, s2, loss, None, grads=grads)
ev[1].record()
torch.cuda.synchronize()
print('ms/step %.4f  loss %.6f' % (ev[0].elapsed_time(ev[1]) / (n - n // 2), float(loss[0] + loss[1])))
