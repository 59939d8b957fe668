"""Time of the tensor-core histogram of listed pairs on the Web-shaped bench batch (back-to-back calls, no L2 flush)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
exec(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), 'web_probe.py')).read().split('for _ in range(2):')[0])
emb, rows = eng.embed_nodes(bd)
for _ in range(5):
    eng.pair_histogram(rows, bd.node_off, g1, g2, bins=16)
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(20):
    eng.pair_histogram(rows, bd.node_off, g1, g2, bins=16)
b.record(); torch.cuda.synchronize()
print('pair_histogram ms', a.elapsed_time(b) / 20)
