import os, sys, time
sys.path.insert(0, '/root/repo')
import numpy as np, torch, bench
from graphembedding_b200.engine import SgnnEngine
from graphembedding_b200.layers_factory import parse_layers
enc, batch, batches = bench.make_train_data(0, 1, n_batches=1)
for name in ('avg_ntn', 'simgnn'):
    spec = bench.TRAIN_MODELS[name][0]
    kw = dict(final_act='sigmoid') if name == 'simgnn' else {}
    eng = SgnnEngine(parse_layers(spec, 29), 29, loss_mode='paired', **kw)
    eng.init_params(123)
    bd = batch.to(eng.device)
    for BIG in (128, 1024, 16384, 131072):
        rng = np.random.default_rng(5)
        b1 = torch.as_tensor(rng.integers(0, bench.TRAIN_GRAPHS, BIG).astype(np.int32), device=eng.device)
        b2 = torch.as_tensor(rng.integers(0, bench.TRAIN_GRAPHS, BIG).astype(np.int32), device=eng.device)
        bt = torch.as_tensor(rng.uniform(0.05, 1, BIG).astype(np.float32), device=eng.device)
        loss = torch.zeros(4, device=eng.device)
        def step():
            eng.train_step(bd, b1, b2, bt, BIG, 1.0, 1.0, loss=loss); eng.apply_adam(1.0, loss)
        for _ in range(5): step()
        torch.cuda.synchronize(); a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(20): step()
        b.record(); torch.cuda.synchronize()
        ms = a.elapsed_time(b) / 20
        print(os.environ.get('SGNN_TRAIN_KERNEL', 'team'), name, BIG, '%.4f ms' % ms, '%.3g pairs/s' % (BIG / ms * 1e3), flush=True)
