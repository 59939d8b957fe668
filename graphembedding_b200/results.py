"""Result objects holding an m x n distance or similarity matrix (reference
``src/results.py``: ``Result`` :10-126, ``DistanceModelResult`` :129-198,
``SimilarityBasedModelResult`` :211-224, ``SiameseModelResult`` :313-347,
``load_result`` :384-395).

Ranking semantics that must not drift:
  * distance results rank ascending with a STABLE sort (ties: lower index first), :143-144;
  * similarity results rank by the stable ascending sort REVERSED (ties: higher index first),
    :221-224;
  * ``top_k_ids(..., inclusive=True)`` extends k over ties of the k-th score, :73-82;
  * ``ranking`` walks back over ties, :103-106.
Tie walks are vectorised here (per-row run boundaries) instead of Python loops."""
import numpy as np

from .similarity import create_sim_kernel


class Result(object):
    def model(self):
        return self.model_

    def m_n(self):
        return self.dist_sim_mat(norm=False).shape

    def dist_sim_mat(self, norm):
        raise NotImplementedError()

    def sort_id_mat(self, norm):
        raise NotImplementedError()

    def _sorted_scores(self, qid, norm):
        ids = self.sort_id_mat(norm)[qid]
        return ids, self.dist_sim_mat(norm)[qid][ids]

    def top_k_ids(self, qid, k, norm, inclusive):
        ids, sc = self._sorted_scores(qid, norm)
        n = len(ids)
        if k < 0 or k >= n:
            raise RuntimeError('Invalid k {}'.format(k))
        if inclusive and k > 0:
            differs = np.nonzero(sc[k:] != sc[k - 1])[0]
            k = k + (int(differs[0]) if len(differs) else n - k)
        return ids[:k]

    def ranking(self, qid, gid, norm, one_based=True):
        ids, sc = self._sorted_scores(qid, norm)
        finds = np.nonzero(ids == gid)[0]
        assert len(finds) == 1
        fid = int(finds[0])
        # ties form one contiguous run in a sorted row: step back to where the run starts
        differs = np.nonzero(sc[:fid] != sc[fid])[0]
        first_equal = int(differs[-1]) + 1 if len(differs) else 0
        return first_equal + 1 if one_based else first_equal

    def time(self, qid, gid):
        raise NotImplementedError()

    def time_mat(self):
        raise NotImplementedError()

    def mat(self, metric, norm):
        raise NotImplementedError()


class DistanceModelResult(Result):
    """Ground-truth style result from an m x n distance matrix.  The reference loads the
    matrix from ``result/<dataset>/<metric>/*.npy`` (:182-189): ``from_disk`` does the same; the
    constructor takes matrices and node counts directly (synthetic ground truth)."""

    def __init__(self, dataset, model, dist_mat, q_nodes, db_nodes, time_mat=None):
        self.dataset, self.model_ = dataset, model
        self.dist_mat_ = np.asarray(dist_mat, np.float64)
        q = np.asarray(q_nodes, np.float64)[:, None]
        d = np.asarray(db_nodes, np.float64)[None, :]
        self.dist_norm_mat_ = 2 * self.dist_mat_ / (q + d)            # distance.py:59-60
        self.time_mat_ = time_mat
        self.sort_id_mat_ = np.argsort(self.dist_mat_, kind='mergesort')
        self.dist_norm_sort_id_mat_ = np.argsort(self.dist_norm_mat_, kind='mergesort')

    def dist_metric(self):
        raise NotImplementedError()

    def dist_mat(self, norm):
        return self.dist_norm_mat_ if norm else self.dist_mat_

    def dist_sim_mat(self, norm):
        return self.dist_mat(norm)

    def dist_sim(self, qid, gid, norm):
        return self.dist_metric(), self.dist_mat(norm)[qid][gid]

    def sim_mat(self, sim_kernel, yeta, norm):
        return create_sim_kernel(sim_kernel, yeta).dist_to_sim_np(self.dist_mat(norm))

    def sort_id_mat(self, norm):
        return self.dist_norm_sort_id_mat_ if norm else self.sort_id_mat_

    def time(self, qid, gid):
        return self.time_mat_[qid][gid]

    def time_mat(self):
        return self.time_mat_

    def mat(self, metric, norm):
        if metric == self.dist_metric():
            return self.dist_mat(norm)
        if metric == 'time':
            return self.time_mat_
        raise RuntimeError('Unknown metric {} for model {}'.format(metric, self.model_))


    # ---- the reference's on-disk format ----
    @classmethod
    def result_root(cls):
        import os
        return os.environ.get('GEB200_RESULT_ROOT', os.path.join(os.path.dirname(os.path.dirname(
            os.path.abspath(__file__))), 'result'))

    @classmethod
    def load_result_mat(cls, dataset, model, dist_metric, metric, root=None):
        """``_load_result_mat`` (src/results.py:182-189): result/<dataset>/<metric>/<dist_metric>_<metric>_mat_
        <dataset>_<model>_*.npy; the first match wins (``_choose_result_file`` :191-192)."""
        from glob import glob
        file_p = (root or cls.result_root()) + '/{}/{}/{}_{}_mat_{}_{}_*.npy'.format(
            dataset, metric, dist_metric, metric, dataset, model)
        li = sorted(glob(file_p))
        if not li:
            raise RuntimeError('No results found {}'.format(file_p))
        return np.load(li[0])

    @classmethod
    def from_disk(cls, dataset, model, root=None):
        """What the reference's constructor does (src/results.py:130-145): distance and time matrices from .npy
        files, node counts for the normalisation (distance.py:59-60) from ``load_data(dataset, True / False)``."""
        from .data import load_data
        probe = cls.__new__(cls)
        dist = cls.load_result_mat(dataset, model, probe.dist_metric(), probe.dist_metric(), root)
        time_mat = cls.load_result_mat(dataset, model, probe.dist_metric(), 'time', root)
        train, test = load_data(dataset, True), load_data(dataset, False)
        m, n = dist.shape
        if m != len(test.graphs) or n != len(train.graphs):
            raise RuntimeError('result matrix {} x {} does not match the dataset ({} test, {} train graphs)'.format(
                m, n, len(test.graphs), len(train.graphs)))
        return cls(dataset, model, dist, [g.number_of_nodes() for g in test.graphs],
                   [g.number_of_nodes() for g in train.graphs], time_mat=time_mat)


class PairwiseGEDModelResult(DistanceModelResult):
    def dist_metric(self):
        return 'ged'


class SimilarityBasedModelResult(Result):
    def sim_mat(self, sim_kernel=None, yeta=None, norm=None):
        return self.sim_mat_

    def dist_sim_mat(self, norm=False):
        return self.sim_mat_

    def dist_sim(self, qid, gid, norm=False):
        return 'sim', self.sim_mat_[qid][gid]

    def sort_id_mat(self, norm=False):
        if getattr(self, 'sort_id_mat_', None) is None:
            self.sort_id_mat_ = np.argsort(self.sim_mat_, kind='mergesort')[:, ::-1]
        return self.sort_id_mat_


class SiameseModelResult(SimilarityBasedModelResult):
    def __init__(self, dataset, model, sim_mat=None, time_mat=None, model_info=None, sort_id_mat=None):
        if sim_mat is None:
            raise NotImplementedError()  # results.py:341-342: loading from disk is not implemented
        self.model_, self.dataset = model, dataset
        self.sim_mat_ = np.asarray(sim_mat)
        self.time_mat_ = time_mat
        # a device-computed ranking (sgnn_sort_rows) may be supplied; it is the same permutation
        self.sort_id_mat_ = None if sort_id_mat is None else np.asarray(sort_id_mat)
        self.sort_id_mat()

    def time(self, qid, gid):
        return self.time_mat_[qid][gid]

    def time_mat(self):
        return self.time_mat_

    def mat(self, metric, *unused):
        if metric == 'sim':
            return self.sim_mat_
        if metric == 'time':
            return self.time_mat_
        raise RuntimeError('Unknown metric {} for model {}'.format(metric, self.model_))


def load_result(dataset, model, sim=None, sim_mat=None, time_mat=None, model_info=None):
    if 'siamese' in model:
        return SiameseModelResult(dataset, model, sim_mat, time_mat, model_info)
    raise RuntimeError('Unknown model {}'.format(model))
