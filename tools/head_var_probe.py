"""Times the SimGNN-style device-resident ranking step (1 000 x 10 000, bf16x3) and prints a hash of the similarity
matrix: run once per SGNN_HEAD_VAR (0 generic epilogue, 1 specialised activations, 2 + constant slot 0, 3 + unrolled
k chunks) to compare the head kernel's epilogue variants -- the hashes must be equal."""
import hashlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from graphembedding_b200.engine import SgnnEngine
from graphembedding_b200.layers_factory import parse_layers
enc, q, db = bench.make_rank_data(0)[:3]
eng = SgnnEngine(parse_layers(bench.SIMGNN_LAYERS, 29), 29, final_act='sigmoid', precision='bf16x3')
eng.init_params(123)
eng.reserve(max_graphs=bench.N_DB, max_nodes=db.total_nodes, max_nnz=db.nnz, max_queries=bench.M_QUERIES, max_db=bench.N_DB)
qd, dbd = q.to(eng.device), db.to(eng.device)
for _ in range(5):
    ids, vals, sim = eng.rank_device(qd, dbd, 10)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
best = 1e9
for rep in range(5):
    a.record()
    for _ in range(40):
        ids, vals, sim = eng.rank_device(qd, dbd, 10)
    b.record()
    torch.cuda.synchronize()
    best = min(best, a.elapsed_time(b) / 40)
h = hashlib.sha1(sim.cpu().numpy().tobytes()).hexdigest()[:16]
print('var', os.environ.get('SGNN_HEAD_VAR', 'default'), 'ms/step %.4f' % best, 'sim sha1', h, 'ids sha1',
      hashlib.sha1(ids.cpu().numpy().tobytes()).hexdigest()[:16])
