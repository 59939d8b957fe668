"""src/distance.py interface.  ``normalized_dist`` (:59-60) is on the hot path; ``ged`` (:23-56)
shells out to an un-vendored Java toolkit in the reference and is out of scope."""


def normalized_dist(d, g1, g2):
    return 2 * d / (g1.number_of_nodes() + g2.number_of_nodes())


def ged(g1, g2, algo, debug=False, timeit=False):
    raise RuntimeError('exact GED needs the Java graph-matching-toolkit, which is not available; '
                       'use dist_calculator.proxy_ged or inject a gidpair_dist_map')
