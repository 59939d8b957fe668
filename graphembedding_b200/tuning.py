"""Hyper-parameter sweep driver with the reference's tables and CSV layout (model/Siamese/tuning.py:19-209;
SURVEY 8f N3).  With a microsecond-scale training step a grid point (500..2000 iterations + the all-pairs test)
costs well under a second, so the reference's 4 800-point grid becomes practical.

Same names as the reference: ``tune``, ``setup_file``, ``parse_results``, ``csv_record``, ``reset_flag``; the
tables ``header``, ``dist_norm_range`` ... ``dropout_range`` are the reference's (checked against
tests/golden/tuning.json, extracted from the reference file).  ``tune`` takes optional ``limit`` / ``grid``
arguments so that tests and quick sweeps can run a slice of the grid; defaults reproduce the full sweep."""
import itertools
import os

import numpy as np

dataset = 'aids80nef'
header = ['dist_norm', 'yeta', 'final_act', 'learning_rate', 'iter', 'validation_ratio',
          'best_train_loss', 'best_train_loss_iter', 'best_val_loss', 'best_val_loss_iter'] + \
         ['prec@{}_norm'.format(i) for i in range(1, 11)] + ['prec@{}_nonorm'.format(i) for i in range(1, 11)] + \
         ['mrr_norm', 'mrr_nonorm', 'mse_norm', 'mse_nonorm']
model = 'siamese_gcntn_mse'
dist_norm_range = [True, False]
yeta_range_norm = [0.1, 0.2, 0.3, 0.4, 0.6, 0.8]
yeta_range_nonorm = [0.01, 0.005, 0.001, 0.0005]
final_act_range = ['identity', 'relu', 'sigmoid', 'tanh', 'sim_kernel']
lr_range = [0.06, 0.03, 0.01, 0.006, 0.003, 0.001]
iter_range = list(range(500, 2001, 500))
val_ratio_range = [0.25, 0.4]
dropout_range = [0.1, 0.5]
layers = [  # tuning.py:74-93
    'GraphConvolution:output_dim=32,act=relu,dropout=True,bias=True,sparse_inputs=True',
    'GraphConvolution:input_dim=32,output_dim=16,act=identity,dropout=True,bias=True,sparse_inputs=False',
    'Average',
    'NTN:input_dim=16,feature_map_dim=10,inneract=relu,dropout=True,bias=True',
]


def grid_points():
    """The sweep in the reference's loop order (tuning.py:54-60)."""
    for dist_norm in dist_norm_range:
        yeta_range = yeta_range_norm if dist_norm else yeta_range_nonorm
        for yeta, final_act, lr, iteration, val_ratio, dropout in itertools.product(
                yeta_range, final_act_range, lr_range, iter_range, val_ratio_range, dropout_range):
            yield dist_norm, yeta, final_act, lr, iteration, val_ratio, dropout


def default_file_name():
    from .utils import get_ts
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    return '{}/logs/parameter_tuning_{}_{}.csv'.format(root, dataset, get_ts())


def setup_file(file_name=None, info=''):
    file_name = file_name or default_file_name()
    os.makedirs(os.path.dirname(os.path.abspath(file_name)), exist_ok=True)
    f = open(file_name, 'w')
    f.write(info)                                   # get_model_info_as_str(), tuning.py:167
    f.write(','.join(map(str, header)) + '\n')
    f.flush()
    return f


def parse_results(results):
    """tuning.py:175-184."""
    mrr_norm = results['mrr_norm'][model]
    mrr_nonorm = results['mrr_nonorm'][model]
    mse_norm = results['mse_norm'][model]
    mse_nonorm = results['mse_nonorm'][model]
    apk_norm = results['apk_norm'][model]['aps'][:10]
    apk_nonorm = results['apk_nonorm'][model]['aps'][:10]
    return list(apk_norm) + list(apk_nonorm) + [mrr_norm, mrr_nonorm, mse_norm, mse_nonorm]


def csv_record(mesg, f):
    f.write(','.join(map(str, mesg)) + '\n')
    f.flush()


def reset_flag(FLAGS, func, name, v):
    """tuning.py:192-194 re-defines the flag; on absl assigning the value is equivalent (``func`` kept for the
    signature)."""
    setattr(FLAGS, name, v)


def tune(FLAGS=None, file_name=None, limit=None, grid=None, run=None):
    """tuning.py:50-158.  ``grid``: iterable of (dist_norm, yeta, final_act, lr, iters, val_ratio, dropout) (default:
    the full reference grid); ``limit``: stop after that many points; ``run``: callable replacing ``train.main``
    (tests).  Returns (best-by-train-loss row, best-by-validation-loss row)."""
    from .config import FLAGS as _FLAGS, ensure_parsed
    FLAGS = FLAGS or _FLAGS
    ensure_parsed()
    if run is None:
        from .train import main as run
    f = setup_file(file_name)
    best_results_train_loss, best_results_val_loss = float('Inf'), float('Inf')
    results_train, results_val = [], []
    for i, (dist_norm, yeta, final_act, lr, iteration, val_ratio, dropout) in enumerate(grid or grid_points(), 1):
        if limit is not None and i > limit:
            break
        print('Number of tuning iteration: {}'.format(i), 'dist_norm: {}'.format(dist_norm), 'yeta: {}'.format(yeta),
              'final_act: {}'.format(final_act), 'learning_rate: {}'.format(lr), 'iteration: {}'.format(iteration),
              'validation_ratio: {}'.format(val_ratio), 'dropout: {}'.format(dropout))
        for name, v in (('dataset', dataset), ('valid_percentage', val_ratio), ('model', model), ('num_layers', 4),
                        ('layer_0', layers[0]), ('layer_1', layers[1]), ('layer_2', layers[2]), ('layer_3', layers[3]),
                        ('dist_norm', dist_norm), ('yeta', yeta), ('final_act', final_act), ('learning_rate', lr),
                        ('iters', iteration), ('dropout', dropout), ('log', False), ('plot_results', False)):
            reset_flag(FLAGS, None, name, v)
        train_costs, train_times, val_costs, val_times, results = run()
        best_train_loss, best_train_iter = np.min(train_costs), np.argmin(train_costs)
        best_val_loss, best_val_iter = np.min(val_costs), np.argmin(val_costs)
        print('best_train_loss: {}'.format(best_train_loss), 'best_val_loss: {}'.format(best_val_loss),
              'best_train_iter: {}'.format(best_train_iter), 'best_val_iter: {}'.format(best_val_iter))
        model_results = parse_results(results)
        csv_record([str(x) for x in [dist_norm, yeta, final_act, lr, iteration, val_ratio, best_train_loss,
                                     best_train_iter, best_val_loss, best_val_iter] + model_results], f)
        if best_train_loss < best_results_train_loss:
            best_results_train_loss = best_train_loss
            results_train = [dist_norm, yeta, final_act, lr, iteration, val_ratio, best_train_loss,
                             best_train_iter] + model_results
        if best_val_loss < best_results_val_loss:
            best_results_val_loss = best_val_loss
            results_val = [dist_norm, yeta, final_act, lr, iteration, val_ratio, best_val_loss,
                           best_val_iter] + model_results
    print(results_train)
    print(results_val)
    tail = [h for h in header if h not in ('best_val_loss', 'best_val_loss_iter')]
    csv_record(['Final results:'], f)
    csv_record(tail, f)
    csv_record([str(x) for x in results_train], f)
    csv_record([h.replace('best_train_loss', 'best_val_loss') for h in tail], f)
    csv_record([str(x) for x in results_val], f)
    f.close()
    return results_train, results_val


if __name__ == '__main__':
    tune()
