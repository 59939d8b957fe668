"""``Eval`` (reference model/Siamese/eval.py:5-42) and the result-dict aggregation of
``src/exp.py`` (``plot_apk`` :337-357 -> ``plot_apk_helper`` :360-369, ``plot_mrr_mse_time``
:422-455) without matplotlib: same dict keys (``apk_norm``, ``apk_nonorm``, ``mrr_*``,
``mse_*``, ``time_*``) and per-model sub-dicts that ``tuning.parse_results`` consumes
(tuning.py:175-184).  ``kendall_tau_*`` keys are additive."""
import numpy as np

from . import metrics
from .results import PairwiseGEDModelResult, load_result


def get_norm_str(norm):
    if norm is None:
        return ''
    return '_norm' if norm else '_nonorm'


class Eval(object):
    def __init__(self, dataset, true_model, sim_kernel_name, yeta, plot_results=False, true_result=None):
        self.dataset = dataset
        self.models = [true_model]
        self.rs = {true_model: true_result}
        self.true_result = true_result
        self.sim_kernel_name = sim_kernel_name
        self.yeta = yeta
        self.results = {}
        self.plot_results = False  # plotting is out of scope

    @staticmethod
    def from_data(data, dist_calculator, model, true_model='astar'):
        """Ground truth from the distance calculator for every (test i, orig-train j) pair
        (the reference loads it from result/<dataset>/ged/*.npy, results.py:182-189)."""
        m, n = data.m_n()
        qs = [data.test_data.get_graph(i).get_nxgraph() for i in range(m)]
        dbs = [data.get_orig_train_graph(j).get_nxgraph() for j in range(n)]
        dist = np.array([[dist_calculator.calculate_dist(q, d)[0] for d in dbs] for q in qs], np.float64)
        tr = PairwiseGEDModelResult(data.dataset, true_model, dist, [g.number_of_nodes() for g in qs],
                                    [g.number_of_nodes() for g in dbs], time_mat=np.zeros((m, n)))
        from .config import FLAGS
        return Eval(data.dataset, true_model, FLAGS.sim_kernel, model.yeta, true_result=tr)

    def get_true_sim(self, query_id, train_id, norm):
        return self.true_result.sim_mat(self.sim_kernel_name, self.yeta, norm)[query_id][train_id]

    def eval_test(self, cur_model, sim_mat, time_mat):
        models = self.models + [cur_model]
        self.rs[cur_model] = load_result(self.dataset, cur_model, sim_mat=sim_mat, time_mat=time_mat)
        _, n = self.true_result.m_n()
        for norm in (True, False):
            ks = range(1, n)                                                   # exp.py:353
            numbers = {mdl: {'ks': ks, 'aps': metrics.precision_at_ks(self.true_result, self.rs[mdl], norm, list(ks))}
                       for mdl in models}
            self.results['apk{}'.format(get_norm_str(norm))] = numbers
        names = ['mrr', 'mse'] + (['time'] if time_mat is not None else [])
        for metric in names:
            for norm in (True, False):
                d = {}
                for mdl in models:
                    if metric == 'mrr':
                        d[mdl] = metrics.mean_reciprocal_rank(self.true_result, self.rs[mdl], norm)
                    elif metric == 'mse':
                        d[mdl] = metrics.mean_squared_error(self.true_result, self.rs[mdl],
                                                            self.sim_kernel_name, self.yeta, norm)
                    else:
                        d[mdl] = metrics.average_time(self.rs[mdl])
                self.results['{}{}'.format(metric, get_norm_str(norm))] = d
        for norm in (True, False):
            self.results['kendall_tau{}'.format(get_norm_str(norm))] = {
                cur_model: metrics.kendall_tau(self.true_result, self.rs[cur_model], norm)[0]}
        return self.results
