import os, sys
sys.path.insert(0, '/root/repo')
import numpy as np, torch, bench
from graphembedding_b200.engine import SgnnEngine
from graphembedding_b200.layers_factory import parse_layers
enc, batch, batches = bench.make_train_data(0, 1, n_batches=1)
eng = SgnnEngine(parse_layers(bench.TRAIN_MODELS['avg_ntn'][0], 29), 29, loss_mode='paired')
eng.init_params(123)
bd = batch.to(eng.device)
g1, g2, t = [torch.as_tensor(x, device=eng.device) for x in batches[0][:3]]
loss = torch.zeros(4, device=eng.device)
for _ in range(6):
    eng.train_step(bd, g1, g2, t, len(g1), 1.0, 1.0, loss=loss); eng.apply_adam(1.0, loss)
torch.cuda.synchronize()
print('ok')
