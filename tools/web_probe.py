"""Web-shaped step (BASELINE configs[4]) a few times, for ncu:
  ncu --set full -k regex:"pair_histogram_tiled|embed_large" -c 4 -o gpurun_out/web python tools/web_probe.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import bench
from graphembedding_b200 import data as D
from graphembedding_b200.engine import SgnnEngine
from graphembedding_b200.graphs import NodeFeatureOneHotEncoder, pack_nx_graphs
from graphembedding_b200.layers_factory import parse_layers
W = type('WebBench', (D.WebData,), {'num_train': 256})
gs = W(True).graphs
enc = NodeFeatureOneHotEncoder(gs, 'type', feat_idx_dic={l: i for i, l in enumerate(D.WebData.labels)})
batch = pack_nx_graphs(gs, enc)
eng = SgnnEngine(parse_layers(bench.TRAIN_LAYERS, 32), 32, precision=os.environ.get('WEB_PRECISION', 'bf16x3'))
eng.init_params(123)
eng.reserve(max_graphs=256, max_nodes=batch.total_nodes, max_nnz=batch.nnz, max_pairs=4096)
bd = batch.to(eng.device)
rng = np.random.default_rng(7)
g1 = torch.as_tensor(rng.integers(0, 256, 4096).astype(np.int32), device=eng.device)
g2 = torch.as_tensor(rng.integers(0, 256, 4096).astype(np.int32), device=eng.device)
for _ in range(2):
    emb, rows = eng.embed_nodes(bd)
    raw, sim = eng.score_pairs(emb, g1, g2)
    hist, counts = eng.pair_histogram(rows, bd.node_off, g1, g2, bins=16)
torch.cuda.synchronize()
print('ok', float(hist.sum()), batch.total_nodes, batch.nnz)
