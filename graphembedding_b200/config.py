"""Flags of the Siamese path: same names, types and defaults as the reference's
``model/Siamese/config.py:4-123`` (which defines them on ``tf.app.flags``, a thin wrapper of
absl), carried here on ``absl.flags`` directly.  Additive flags (not in the reference) are at
the bottom."""
from absl import flags

_D = {}


def _def(kind, name, default, help_):
    if name in flags.FLAGS:
        return
    getattr(flags, 'DEFINE_' + kind)(name, default, help_)
    _D[name] = default


# data
_def('string', 'dataset', 'aids80nef', 'Dataset string.')
_def('float', 'valid_percentage', 0.25, '(# validation graphs) / (# validation + # training graphs.')
_def('string', 'node_feat_name', 'type', 'Name of the node feature.')
_def('string', 'node_feat_encoder', 'onehot', 'How to encode the node feature.')
_def('string', 'edge_feat_name', 'valence', 'Name of the edge feature.')
_def('string', 'edge_feat_processor', None, 'How to process the edge feature.')
_def('string', 'dist_metric', 'ged', 'Distance metric to use.')
_def('string', 'dist_algo', 'astar', 'Ground-truth distance algorithm to use.')
_def('string', 'sampler', 'random', 'Sampler to use.')
_def('integer', 'sample_num', -1, 'Number of pairs to sample for training.')
_def('boolean', 'sampler_duplicate_removal', False, 'Whether to remove duplicate for sampler or not.')
# model
_def('string', 'model', 'siamese_gcntn_mse', 'Model string.')
_def('integer', 'num_layers', 5, 'Number of layers.')
_def('string', 'layer_0',
     'GraphConvolution:output_dim=32,act=relu,dropout=True,bias=True,sparse_inputs=True', '')
_def('string', 'layer_1',
     'GraphConvolution:input_dim=32,output_dim=16,act=identity,dropout=True,bias=True,sparse_inputs=False', '')
_def('string', 'layer_2', 'Dense:input_dim=16,output_dim=1,dropout=True,act=relu,bias=True', '')
_def('string', 'layer_3', 'Padding:max_in_dims=10,padding_value=0', '')
_def('string', 'layer_4', 'NTN:input_dim=10,feature_map_dim=10,inneract=relu,dropout=True,bias=True', '')
for _i in range(5, 12):  # room for deeper stacks (the reference re-defines flags in tuning.py:192-194)
    _def('string', 'layer_%d' % _i, None, '')
_def('integer', 'batch_size', 5, 'Number of graph pairs in a batch.')
_def('boolean', 'dist_norm', True, 'Whether to normalize the distance or not '
     'when choosing the ground truth distance.')
_def('string', 'sim_kernel', 'gaussian', 'Name of the similarity kernel.')
_def('float', 'yeta', 0.6, 'yeta for the gaussian kernel function.')
_def('string', 'final_act', 'sim_kernel', 'The final activation function applied to the NTN output.')
_def('string', 'loss_func', 'mse', 'Loss function(s) to use.')
_def('float', 'delta', 0.1, 'Margin between positive pairs ground truth scores'
     'and negative pairs scores  ground truth scores')
_def('float', 'gamma', 0.1, 'Margin between positive pairs prediction scores'
     'and negative pairs prediction scores')
_def('integer', 'num_neg', 5, 'Number of negative samples.')
_def('string', 'graph_loss', None, 'Loss function(s) to use.')
_def('float', 'dropout', 0.1, 'Dropout rate (1 - keep probability).')
_def('float', 'weight_decay', 5e-4, 'Weight for L2 loss on embedding matrix.')
_def('float', 'learning_rate', 0.01, 'Initial learning rate.')
# train / validate / test
_def('integer', 'iters', 20, 'Number of iterations to train.')
_def('integer', 'early_stopping', None, 'Tolerance for early stopping (# of iters).')
_def('boolean', 'log', True, 'Whether to log the results via Tensorboard or not.')
_def('boolean', 'plot_results', False, 'Whether to plot the results (involving all baselines) or not.')

# ---- additive flags (not in the reference) ----
_def('string', 'ntn_mode', 'reference', "'reference' reproduces the (K,1)*(K,1,1) broadcast of "
     "layers.py:305-308 (score = sum(U) * sum(act(m))); 'paper' is sum_k U_k act(m_k).")
_def('string', 'loss_mode', 'reference', "'reference' reproduces the (B,1)-(B,) broadcast of "
     "model_mse.py:145-151; 'paired' is the per-pair MSE.")
_def('string', 'label_mode', 'reference', "'reference' draws the labels from B fresh pairs as "
     "model_mse.py:82-92 does (F7); 'paired' labels the pairs that are fed.")
_def('string', 'precision', 'fp32', 'all-pairs scorer arithmetic: fp32 | bf16x3 | bf16')
_def('integer', 'device', 0, 'CUDA device ordinal')
_def('integer', 'seed', 123, 'seed of the glorot initialiser')

FLAGS = flags.FLAGS


def ensure_parsed(argv=None):
    """Mark flags as parsed with defaults when the host program has not parsed argv."""
    if not FLAGS.is_parsed():
        FLAGS(argv or ['graphembedding_b200'])
    return FLAGS


def layer_specs():
    d = FLAGS.flag_values_dict()
    return [d['layer_{}'.format(i)] for i in range(FLAGS.num_layers)]
