"""Host pre-processing of graphs for the device path.

Mirror of the reference's ``model/Siamese/graphs.py`` (``ModelGraphList`` :8-31,
``ModelGraph`` :34-95, ``NodeFeatureOneHotEncoder`` :98-117): same classes and accessors,
but a whole list of graphs is normalised in ONE call of the native packer
(``sgnn_pack_graphs``) into the block-diagonal CSR batch the kernels consume, instead of one
scipy pipeline per graph.
"""
import networkx as nx
import numpy as np

from .native import PackedBatch
from .samplers import DistributionSampler, RandomSampler


class NodeFeatureOneHotEncoder(object):
    """Label string -> one-hot column.  The reference enumerates a Python ``set``
    (graphs.py:101-106), so its column order depends on the process' string hash seed (F11);
    here labels are sorted for a reproducible order unless an explicit ``feat_idx_dic`` (e.g.
    taken from a reference encoder object) is supplied."""

    def __init__(self, gs, node_feat_name, feat_idx_dic=None):
        self.node_feat_name = node_feat_name
        if feat_idx_dic is None:
            feats = set()
            for g in gs:
                feats |= set(nx.get_node_attributes(g, node_feat_name).values())
            feat_idx_dic = {f: i for i, f in enumerate(sorted(feats, key=str))}
        self.feat_idx_dic = dict(feat_idx_dic)

    def labels(self, g):
        """Integer one-hot column of every node, in ``g.nodes()`` order (graphs.py:108-110)."""
        attr = nx.get_node_attributes(g, self.node_feat_name)
        return np.array([self.feat_idx_dic[attr[n]] for n in g.nodes()], np.int32)

    def encode(self, g):
        lab = self.labels(g)
        out = np.zeros((len(lab), self.input_dim()))
        out[np.arange(len(lab)), lab] = 1.0
        return out

    def input_dim(self):
        return len(self.feat_idx_dic)


def _edge_index(g):
    pos = {n: i for i, n in enumerate(g.nodes())}
    return np.array([(pos[u], pos[v]) for u, v in g.edges()], np.int32).reshape(-1, 2)


def pack_nx_graphs(gs, encoder):
    """networkx graphs -> PackedBatch (host numpy arrays)."""
    return PackedBatch.from_lists([encoder.labels(g) for g in gs], [_edge_index(g) for g in gs])


class ModelGraph(object):
    """One graph of a packed batch (reference graphs.py:34-53 accessors)."""

    def __init__(self, nxgraph, node_feat_encoder, _batch=None, _index=0):
        self.nxgraph = nxgraph
        if _batch is None:
            _batch = pack_nx_graphs([nxgraph], node_feat_encoder)
            _index = 0
        self.batch, self.index = _batch, _index
        self._input_dim = node_feat_encoder.input_dim()

    def get_nxgraph(self):
        return self.nxgraph

    def num_nodes(self):
        b, g = self.batch, self.index
        return int(b.node_off[g + 1] - b.node_off[g])

    def labels(self):
        b, g = self.batch, self.index
        return np.asarray(b.labels[b.node_off[g]:b.node_off[g + 1]])

    def csr(self):
        """(rowptr, col, val) of A_hat local to this graph."""
        b, g = self.batch, self.index
        lo, hi = int(b.node_off[g]), int(b.node_off[g + 1])
        rp = np.asarray(b.rowptr[lo:hi + 1])
        return (rp - rp[0]).astype(np.int32), np.asarray(b.col[rp[0]:rp[-1]]), np.asarray(b.val[rp[0]:rp[-1]])

    def get_node_inputs(self):
        """Row-normalised one-hot features as the reference's COO tuple (graphs.py:55-62,
        78-95): (coords int32 [N,2], values float64 [N], shape)."""
        lab = self.labels()
        n = len(lab)
        coords = np.stack([np.arange(n, dtype=np.int32), lab.astype(np.int32)], 1)
        return coords, np.ones(n, np.float64), (n, self._input_dim)

    def get_node_inputs_num_nonzero(self):
        return self.get_node_inputs()[1].shape

    def get_laplacians(self):
        """[A_hat] as the reference's COO tuple, in the reference's entry order (sorted by
        column, then row: ``.transpose()...tocoo()``, graphs.py:76, 81-85).  Values are the
        float32 the device consumes, widened to float64."""
        rp, col, val = self.csr()
        n = len(rp) - 1
        rows = np.repeat(np.arange(n, dtype=np.int32), np.diff(rp))
        coords = np.stack([col.astype(np.int32), rows], 1)  # symmetric: (j, i) listing
        return [(coords, val.astype(np.float64), (n, n))]


class ModelGraphList(object):
    """Reference graphs.py:8-31."""

    def __init__(self, sampler, sample_num, sampler_duplicate_removal, gs, node_feat_encoder):
        self.batch = pack_nx_graphs(gs, node_feat_encoder)
        self.gs = [ModelGraph(g, node_feat_encoder, self.batch, i) for i, g in enumerate(gs)]
        if sampler == 'random':
            self.sampler = RandomSampler(self.gs, sample_num, sampler_duplicate_removal)
        elif sampler == 'density':
            self.sampler = DistributionSampler(self.gs, sample_num, sampler_duplicate_removal)
        else:
            raise RuntimeError('Unknown sampler {}'.format(sampler))

    def num_graphs(self):
        return len(self.gs)

    def get_graph_pair(self):
        return self.sampler.get_pair()

    def get_triple_for_hinge_loss(self):
        return self.sampler.get_triple_for_hinge_loss()

    def get_graph(self, id):
        return self.gs[id]
