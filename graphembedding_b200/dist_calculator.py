"""Ground-truth distance lookup behind the reference's ``DistCalculator.calculate_dist``
interface (model/Siamese/dist_calculator.py:26-44): returns ``(d, normalized_d)``.

The reference shells out to a Java graph-matching toolkit for exact GED
(src/distance.py:23-56); that toolkit is not vendored and Java is unavailable, so the
backing distance here is a deterministic integer proxy (a lower bound of unit-cost GED:
label-multiset mismatch + edge-count difference).  A real ``(gid, gid) -> d`` table can be
injected with ``gidpair_dist_map``."""
from collections import Counter, OrderedDict

from .distance import normalized_dist


def proxy_ged(g1, g2, node_feat_name='type'):
    c1 = Counter(d.get(node_feat_name) for _, d in g1.nodes(data=True))
    c2 = Counter(d.get(node_feat_name) for _, d in g2.nodes(data=True))
    common = sum((c1 & c2).values())
    return int(max(g1.number_of_nodes(), g2.number_of_nodes()) - common +
               abs(g1.number_of_edges() - g2.number_of_edges()))


class DistCalculator(object):
    def __init__(self, dataset, dist_metric, algo, gidpair_dist_map=None, dist_func=None):
        if dist_metric != 'ged':
            raise RuntimeError('Unknwon distance metric {}'.format(dist_metric))
        self.dataset, self.algo = dataset, algo
        self.gidpair_dist_map = OrderedDict(gidpair_dist_map or {})
        self.dist_func = dist_func or (lambda g1, g2, algo: proxy_ged(g1, g2))

    def calculate_dist(self, g1, g2):
        pair = (g1.graph['gid'], g2.graph['gid'])
        d = self.gidpair_dist_map.get(pair)
        if d is None:
            rev = self.gidpair_dist_map.get((pair[1], pair[0]))
            d = rev if rev else self.dist_func(g1, g2, self.algo)  # dist_calculator.py:33-37
            self.gidpair_dist_map[pair] = d
        return d, normalized_dist(d, g1, g2)
