"""Random-shape sweep of the SimGNN-style ranking step on the tensor-core path (histogram GEMM + head kernel with the
second MMA + slice hints + hinted top-k) against the CUDA-core fp32 engine and against its own full matrix."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tests'))
import numpy as np, torch
from util import CONFIGS, O, corpus, encoder_for
from graphembedding_b200.engine import SgnnEngine
from graphembedding_b200.graphs import pack_nx_graphs
from graphembedding_b200.layers_factory import parse_layers

rng = np.random.default_rng(int(os.environ.get('FUZZ_SEED', '1')))
layers = parse_layers(CONFIGS['simgnn'], 29)
flat = O.init_params(O.parse_layer_specs(CONFIGS['simgnn'], 29), seed=5)
flat = flat + (np.random.default_rng(9).standard_normal(flat.shape) * 0.01).astype(np.float32)
engs = {}
for prec in ('bf16x3', 'fp32'):
    e = SgnnEngine(layers, 29, precision=prec, final_act='sigmoid')
    e.set_params(flat)
    engs[prec] = e
gs = corpus(5000, 1, 10, 17)
enc = encoder_for(gs)
bad = 0
for case in range(int(os.environ.get('FUZZ_CASES', '24'))):
    m = int(rng.choice([1, 2, 7, 8, 9, 15, 16, 17, 100, 129])); n = int(rng.choice([1, 31, 127, 128, 129, 511, 512, 513, 1024, 1500, 4097]))
    k = int(min(n, rng.choice([1, 5, 10, 32])))
    qi = rng.choice(len(gs), m, replace=False); di = rng.choice(len(gs), n, replace=False)
    q, db = pack_nx_graphs([gs[i] for i in qi], enc), pack_nx_graphs([gs[i] for i in di], enc)
    out = {}
    for prec, e in engs.items():
        e.reserve(max_graphs=n, max_nodes=max(q.total_nodes, db.total_nodes), max_nnz=db.nnz, max_queries=m, max_db=n)
        for rep in range(3):      # eager, capture, replay
            ids, vals, sim = e.rank_device(q.to(e.device), db.to(e.device), k)
        torch.cuda.synchronize()
        ids2, vals2 = e.topk(sim.contiguous(), k)
        torch.cuda.synchronize()
        out[prec] = (ids.cpu().numpy(), sim.cpu().numpy(), bool(torch.equal(ids, ids2) and torch.equal(vals, vals2)))
    diff = np.abs(out['bf16x3'][1] - out['fp32'][1])
    frac = float((diff > 1e-4).mean())
    good = out['bf16x3'][2] and out['fp32'][2] and frac <= 0.02 and float(diff.max()) < 0.2
    bad += not good
    print('rank', case, 'm', m, 'n', n, 'k', k, 'ok' if good else 'FAIL', out['bf16x3'][2], frac, float(diff.max()), flush=True)
print('RANK FUZZ', 'OK' if bad == 0 else 'FAILED %d' % bad)
