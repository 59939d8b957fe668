"""Dataset objects for the Siamese path.

Drop-in for the reference's ``src/data.py`` (classes ``Data`` :9-33, ``SynData``
:36-59, ``AIDSData`` :62-83, ``AIDS10kData`` :95-100, ``AIDS10kNEFData`` :103-109,
``AIDS700nefData`` :112-121, ``AIDS80nefData`` :124-132): every object exposes
``.graphs`` -- a list of ``networkx.Graph`` with ``graph['gid']`` and a node
attribute ``type`` -- plus ``get_gids()``.

The real AIDS / Web corpora are not available offline, so each loader first
looks for ``<data_root>/<Folder>/{train,test}/*.gexf`` (the reference's on-disk
layout, ``data.py:69-75``) and otherwise builds a seeded synthetic corpus of
the same shape.  ``Web`` has no loader in the reference
(``src/data_preprocess/preprocess_web.py`` is empty); the synthetic ``WebData``
here is build-defined.
"""
import os
import random
import re
from glob import glob

import networkx as nx
import numpy as np

# 29 node labels, carbon/oxygen/nitrogen-heavy like AIDS (BASELINE.json: "29 one-hot labels")
AIDS_LABELS = ['C', 'O', 'N', 'Cl', 'F', 'S', 'Se', 'P', 'Na', 'I', 'Co', 'Br', 'Li', 'Si', 'Mg',
               'Cu', 'As', 'B', 'Pt', 'Ru', 'K', 'Pd', 'Au', 'Te', 'W', 'Rh', 'Zn', 'Bi', 'Pb']
assert len(AIDS_LABELS) == 29

_DATA_ROOT = os.environ.get('GEB200_DATA_ROOT', os.path.join(os.path.dirname(os.path.dirname(
    os.path.abspath(__file__))), 'data'))


def _label_probs(num_labels):
    # Zipf-like: p_i ~ 1/(i+1)^1.6 -> C ~ 55 %, O ~ 18 %, N ~ 10 %, long tail
    w = 1.0 / np.power(np.arange(1, num_labels + 1, dtype=np.float64), 1.6)
    return w / w.sum()


def synthetic_molecule(rng, n, labels, probs, extra_edge_rate=0.12, gid=0):
    """Connected sparse graph on ``n`` nodes: random spanning tree with a degree
    cap of 4 plus a few ring-closing edges (E ~ N, molecule-like)."""
    g = nx.Graph()
    lab = rng.choice(len(labels), size=n, p=probs)
    for i in range(n):
        g.add_node(str(i), type=labels[lab[i]])
    deg = np.zeros(n, np.int64)
    for i in range(1, n):
        cand = [j for j in range(i) if deg[j] < 4]
        j = cand[int(rng.integers(len(cand)))] if cand else int(rng.integers(i))
        g.add_edge(str(j), str(i))
        deg[i] += 1
        deg[j] += 1
    n_extra = int(rng.binomial(max(n - 2, 0), extra_edge_rate))
    for _ in range(n_extra):
        u, v = int(rng.integers(n)), int(rng.integers(n))
        if u != v and not g.has_edge(str(u), str(v)) and deg[u] < 4 and deg[v] < 4:
            g.add_edge(str(u), str(v))
            deg[u] += 1
            deg[v] += 1
    g.graph['gid'] = gid
    return g


def synthetic_corpus(num, n_lo, n_hi, seed, gid_base=0, labels=AIDS_LABELS, size_skew=True,
                     extra_edge_rate=0.12):
    """``num`` connected graphs with node counts in [n_lo, n_hi]."""
    rng = np.random.default_rng(seed)
    probs = _label_probs(len(labels))
    sizes = np.arange(n_lo, n_hi + 1)
    if size_skew:  # AIDS700nef is dominated by the largest sizes (mean ~ 8.9 of max 10)
        w = np.power(np.linspace(0.15, 1.0, len(sizes)), 3.0)
        w = w / w.sum()
    else:
        w = np.full(len(sizes), 1.0 / len(sizes))
    out = []
    for i in range(num):
        n = int(rng.choice(sizes, p=w))
        out.append(synthetic_molecule(rng, n, labels, probs, extra_edge_rate, gid=gid_base + i))
    return out


def _nat_key(s):
    return [int(t) if t.isdigit() else t for t in re.split('([0-9]+)', s)]


class Data(object):
    """Base: ``.graphs`` + ``get_gids`` (reference ``data.py:9-33``; the
    reference's pickle cache in ``save/`` is an on-disk format)."""

    def __init__(self, name_str=''):
        self.name = self.__class__.__name__ + '_' + name_str + self.name_suffix()
        self.init()

    def init(self):
        raise NotImplementedError()

    def name_suffix(self):
        return ''

    def get_gids(self):
        return [g.graph['gid'] for g in self.graphs]


class SynData(Data):
    """Random G(n, m) graphs (reference ``data.py:36-59``)."""
    train_num_graphs = 20
    test_num_graphs = 10

    def __init__(self, train):
        self.num_graphs = SynData.train_num_graphs if train else SynData.test_num_graphs
        self.train = train
        super().__init__('train' if train else 'test')

    def init(self):
        r = random.Random(123 if self.train else 321)
        self.graphs = []
        for i in range(self.num_graphs):
            n = r.randint(5, 20)
            m = r.randint(n - 1, n * (n - 1) // 2)
            g = nx.gnm_random_graph(n, m, seed=r.randint(0, 1 << 30))
            g.graph['gid'] = i
            self.graphs.append(g)

    def name_suffix(self):
        return '_{}_{}'.format(SynData.train_num_graphs, SynData.test_num_graphs)


class AIDSData(Data):
    folder = None
    strip_valence = False   # the *nef variants drop the edge attribute (no edge features)
    synthetic_shape = None  # (train_num, test_num, n_lo, n_hi)

    def __init__(self, train):
        self.train = train
        super().__init__('train' if train else 'test')

    def init(self):
        d = '{}/{}/{}'.format(_DATA_ROOT, self.folder, 'train' if self.train else 'test')
        files = sorted(glob(d + '/*.gexf'), key=_nat_key)
        self.synthetic = not files
        if files:  # reference on-disk layout, data.py:69-78
            self.graphs = []
            for f in files:
                g = nx.read_gexf(f)
                g.graph['gid'] = int(os.path.basename(f).split('.')[0])
                if not nx.is_connected(g):
                    raise RuntimeError('{} not connected'.format(g.graph['gid']))
                self.graphs.append(g)
        else:
            import warnings
            warnings.warn('{}: no .gexf files under {} -- using the SYNTHETIC {}-shaped corpus (set GEB200_DATA_ROOT to '
                          'the reference data/ directory for real graphs)'.format(type(self).__name__, d, self.folder))
            self.graphs = self._synthesize()
        if self.strip_valence or 'nef' in self.folder:  # data.py:80-83, 119-121
            for g in self.graphs:
                for _, _, ed in g.edges(data=True):
                    ed.pop('valence', None)

    def _synthesize(self):
        ntr, nte, lo, hi = self.synthetic_shape
        if self.train:
            return synthetic_corpus(ntr, lo, hi, seed=123, gid_base=0)
        return synthetic_corpus(nte, lo, hi, seed=124, gid_base=ntr)


class AIDS700nefData(AIDSData):
    """560 train / 140 test, <= 10 nodes (``preprocess_aids.py:139-143``)."""
    folder = 'AIDS700nef'
    synthetic_shape = (560, 140, 2, 10)


class AIDS80nefData(AIDS700nefData):
    """Seeded shuffle of AIDS700nef, first 70 / 10 (reference ``data.py:124-132``)."""

    def init(self):
        self.graphs = AIDS700nefData(self.train).graphs
        random.Random(123).shuffle(self.graphs)
        self.graphs = self.graphs[0:70] if self.train else self.graphs[0:10]


class AIDS10kData(AIDSData):
    """10 000 train graphs with <= 30 nodes; test = one graph per size 5..22
    (``preprocess_aids.py:130-138``)."""
    folder = 'AIDS10k'
    synthetic_shape = (10000, 18, 2, 30)

    def _synthesize(self):
        if self.train:
            return synthetic_corpus(10000, 2, 30, seed=223, gid_base=0, size_skew=False)
        rng = np.random.default_rng(224)
        probs = _label_probs(len(AIDS_LABELS))
        return [synthetic_molecule(rng, n, AIDS_LABELS, probs, gid=10000 + i)
                for i, n in enumerate(range(5, 23))]


class AIDS10kNEFData(AIDS10kData):
    """Loads AIDS10k and strips the valence edge attribute (reference ``data.py:95-109``: there is no separate folder)."""
    folder = 'AIDS10k'
    strip_valence = True


class RankingData(AIDSData):
    """Config 4: 1 000 query x 10 000 database AIDS-shaped graphs (<= 10 nodes)."""
    folder = 'AIDSrank'
    synthetic_shape = (10000, 1000, 2, 10)


class WebData(Data):
    """Web-shaped: N ~ U[100, 500], E ~ 2N, connected, sparse (build-defined, F12)."""
    num_train, num_test = 64, 16
    labels = ['w{}'.format(i) for i in range(32)]

    def __init__(self, train):
        self.train = train
        super().__init__('train' if train else 'test')

    def init(self):
        rng = np.random.default_rng(323 if self.train else 324)
        probs = _label_probs(len(self.labels))
        num = self.num_train if self.train else self.num_test
        base = 0 if self.train else self.num_train
        self.graphs = []
        for i in range(num):
            n = int(rng.integers(100, 501))
            g = nx.Graph()
            lab = rng.choice(len(self.labels), size=n, p=probs)
            for v in range(n):
                g.add_node(str(v), type=self.labels[lab[v]])
            for v in range(1, n):  # spanning tree
                g.add_edge(str(int(rng.integers(v))), str(v))
            extra = rng.integers(0, n, size=(n, 2))  # ~N more edges -> E ~ 2N
            for u, v in extra:
                if u != v:
                    g.add_edge(str(int(u)), str(int(v)))
            g.graph['gid'] = base + i
            self.graphs.append(g)


def load_data(data, train):
    """Reference ``src/utils.py:15-35`` (same dataset names; ``aidsrank`` and
    ``web`` are additions)."""
    table = {'syn': SynData, 'aids10k': AIDS10kData, 'aids10knef': AIDS10kNEFData,
             'aids700nef': AIDS700nefData, 'aids80nef': AIDS80nefData,
             'aidsrank': RankingData, 'web': WebData}
    if data not in table:
        raise RuntimeError('Not recognized data %s' % data)
    return table[data](train)
