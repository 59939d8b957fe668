"""Shared helpers for the tests: build the same inputs for the CUDA path and the oracle."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import siamese_oracle as O  # noqa: E402  (tests are allowed to import the oracle)
from graphembedding_b200 import data as D  # noqa: E402
from graphembedding_b200.graphs import NodeFeatureOneHotEncoder, pack_nx_graphs, _edge_index  # noqa: E402

DEFAULT = list(O.DEFAULT_LAYERS)
AVG_NTN = [  # tuning.py:74-93 shape: GCN, GCN, Average, NTN(16, K=10)
    'GraphConvolution:output_dim=32,act=relu,dropout=True,bias=True,sparse_inputs=True',
    'GraphConvolution:input_dim=32,output_dim=16,act=identity,dropout=True,bias=True,sparse_inputs=False',
    'Average',
    'NTN:input_dim=16,feature_map_dim=10,inneract=relu,dropout=True,bias=True',
]
ATT3 = [  # SimGNN-style node stack: 3 GCN + Attention + NTN(16,16)
    'GraphConvolution:output_dim=64,act=relu,dropout=True,bias=True,sparse_inputs=True',
    'GraphConvolution:input_dim=64,output_dim=32,act=relu,dropout=True,bias=True,sparse_inputs=False',
    'GraphConvolution:input_dim=32,output_dim=16,act=tanh,dropout=True,bias=True,sparse_inputs=False',
    'Attention:input_dim=16',
    'NTN:input_dim=16,feature_map_dim=16,inneract=tanh,dropout=True,bias=True',
]
DENSE_DOT = [
    'GraphConvolution:output_dim=24,act=sigmoid,dropout=False,bias=False,sparse_inputs=True',
    'Dense:input_dim=24,output_dim=12,dropout=False,act=tanh,bias=True',
    'GraphConvolution:input_dim=12,output_dim=8,act=relu,dropout=False,bias=True,sparse_inputs=False',
    'Average',
    'Dot',
]
SIMGNN = list(O.SIMGNN_LAYERS)   # BASELINE configs[0]: 3 GCN + Attention + NTN(16,16) + Histogram(16) + FC 32-16-8-4-1
HEAD_NOHIST = AVG_NTN + [        # Dense head straight on the NTN's K-vector
    'Dense:input_dim=10,output_dim=4,dropout=False,act=tanh,bias=True',
    'Dense:input_dim=4,output_dim=1,dropout=False,act=identity,bias=False',
]
HEAD_DROP = [                    # head with dropout on its Dense inputs, odd widths, 7 bins
    'GraphConvolution:output_dim=20,act=relu,dropout=True,bias=True,sparse_inputs=True',
    'GraphConvolution:input_dim=20,output_dim=12,act=tanh,dropout=True,bias=True,sparse_inputs=False',
    'Average',
    'NTN:input_dim=12,feature_map_dim=6,inneract=sigmoid,dropout=True,bias=True',
    'Histogram:bins=7',
    'Dense:input_dim=13,output_dim=9,dropout=True,act=sigmoid,bias=True',
    'Dense:input_dim=9,output_dim=5,dropout=True,act=relu,bias=True',
    'Dense:input_dim=5,output_dim=1,dropout=False,act=identity,bias=True',
]
CONFIGS = {'simgnn': SIMGNN, 'head_nohist': HEAD_NOHIST, 'head_drop': HEAD_DROP, 'default': DEFAULT, 'avg_ntn': AVG_NTN, 'att3': ATT3, 'dense_dot': DENSE_DOT}


def corpus(num, n_lo=2, n_hi=10, seed=7):
    return D.synthetic_corpus(num, n_lo, n_hi, seed)


def encoder_for(gs):
    # fixed 29-label vocabulary so that input_dim == 29 regardless of which labels occur
    return NodeFeatureOneHotEncoder(gs, 'type', feat_idx_dic={l: i for i, l in enumerate(D.AIDS_LABELS)})


def oracle_graphs(gs, enc):
    return [O.Graph.from_edges(enc.labels(g), [tuple(e) for e in _edge_index(g)]) for g in gs]


def synthetic_ged(g1, g2):
    from graphembedding_b200.dist_calculator import proxy_ged
    return proxy_ged(g1, g2)


def rel_err(a, b, floor=1e-6):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return np.max(np.abs(a - b) / np.maximum(np.abs(b), floor)) if a.size else 0.0


# ---- tolerances (ONE number per dtype) ----
# fp32 paths (CUDA-core kernels AND the bf16x3 tensor-core kernels, which are fp32-grade): raw scores and
# similarities within TOL_FP32 * max(1, max |raw score|) ABSOLUTE of the fp64 oracle.  (A similarity is
# final_act(raw) with |d final_act / d raw| <= 0.66 for exp(-0.6 x^2) and <= 0.25 for the sigmoid, so the raw-score
# gate implies the similarity gate; relative gates on values near 0 are not used.)
TOL_FP32 = 1e-4
TOL_GRAD = 2e-4          # gradients: fraction of max |g| (fp32 atomics accumulate B pairs in arbitrary order)
TOL_BF16 = 1e-1          # plain bf16 operands (one hi.hi product): relative, floor 1e-2


def close_fp32(got, ref, raw_scale=1.0):
    got, ref = np.asarray(got, np.float64), np.asarray(ref, np.float64)
    return got.size == 0 or np.max(np.abs(got - ref)) <= TOL_FP32 * max(1.0, float(raw_scale))
