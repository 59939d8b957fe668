"""Layer-spec grammar of the reference (model/Siamese/layers_factory.py:8-135):
``'Name:key=val,key=val'`` strings -> plain dict descriptions that ``native.make_config``
turns into the POD ``SgnnLayerDesc[]`` handed to ``sgnn_create``.  Same layer names, same
arity checks and error behaviour (RuntimeError); ``Attention`` (layers.py:143-160, defined
but never wired into the reference's factory) is accepted as an addition, and so is the
SimGNN-style head (build-defined): ``Histogram:bins=B`` right after the NTN and ``Dense``
layers after that, the last with ``output_dim=1`` (include/sgnn.h, SGNN_LAYER_HISTOGRAM)."""
import numpy as np

_ARITY = {'GraphConvolution': (5, 6), 'Average': (0, 0), 'NTN': (5, 5), 'Dot': (0, 0),
          'Dense': (5, 5), 'Padding': (2, 2), 'Attention': (1, 1), 'Histogram': (1, 1)}
_ARITY_MSG = {'GraphConvolution': 'GraphConvolution layer must have 3-4 specs',
              'Average': 'Average layer must have 0 specs',
              'NTN': 'Average layer must have 0 specs',   # sic: reference message, :67
              'Dot': 'Dot layer must have 0 specs',
              'Dense': 'Dot layer must have 5 specs',     # sic: reference message, :84
              'Padding': 'Padding layer must have 2 specs',
              'Attention': 'Attention layer must have 1 spec',
              'Histogram': 'Histogram layer must have 1 spec'}
ACTIVATIONS = ('relu', 'identity', 'sigmoid', 'tanh', 'sim_kernel')


def parse_as_bool(b):
    if b == 'True':
        return True
    if b == 'False':
        return False
    raise RuntimeError('Unknown bool string {}'.format(b))


def _check_act(a):
    if a not in ACTIVATIONS:
        raise RuntimeError('Unknown activation function {}'.format(a))
    return a


def parse_layer(spec, layer_id, model_input_dim):
    head, _, tail = spec.partition(':')
    if ':' in tail:
        raise AssertionError('layer spec has more than one ":"')
    kv = {}
    if tail:
        for item in tail.split(','):
            k, v = item.split('=')
            kv[k] = v
    if head not in _ARITY:
        raise RuntimeError('Unknown layer {}'.format(head))
    lo, hi = _ARITY[head]
    if not lo <= len(kv) <= hi:
        raise RuntimeError(_ARITY_MSG[head])
    out = {'kind': head}
    if head == 'GraphConvolution':
        ind = kv.get('input_dim')
        if not ind:
            if layer_id != 0:
                raise RuntimeError('The input dim for layer {} must be specified'.format(layer_id))
            ind = model_input_dim
        out.update(input_dim=int(ind), output_dim=int(kv['output_dim']),
                   dropout=parse_as_bool(kv['dropout']), sparse_inputs=parse_as_bool(kv['sparse_inputs']),
                   act=_check_act(kv['act']), bias=parse_as_bool(kv['bias']))
    elif head == 'Dense':
        out.update(input_dim=int(kv['input_dim']), output_dim=int(kv['output_dim']),
                   dropout=parse_as_bool(kv['dropout']), act=_check_act(kv['act']),
                   bias=parse_as_bool(kv['bias']))
    elif head == 'NTN':
        out.update(input_dim=int(kv['input_dim']), feature_map_dim=int(kv['feature_map_dim']),
                   dropout=parse_as_bool(kv['dropout']), inneract=_check_act(kv['inneract']),
                   bias=parse_as_bool(kv['bias']))
    elif head == 'Padding':
        out.update(max_in_dims=int(kv['max_in_dims']), padding_value=int(kv['padding_value']))
    elif head == 'Attention':
        out.update(input_dim=int(kv['input_dim']))
    elif head == 'Histogram':
        out.update(bins=int(kv['bins']))
    return out


def parse_layers(spec_strings, model_input_dim):
    return [parse_layer(s, i, model_input_dim) for i, s in enumerate(spec_strings)]


def create_layers(model):
    """Reference signature (layers_factory.py:8): reads ``layer_i`` from FLAGS."""
    from .config import FLAGS
    d = FLAGS.flag_values_dict()
    return parse_layers([d['layer_{}'.format(i)] for i in range(FLAGS.num_layers)], model.input_dim)


# numpy activation twins (layers_factory.py:117-126), used on host-side results
def relu_np(x):
    return np.maximum(x, 0)


def identity_np(x):
    return x


def sigmoid_np(x):
    return 1 / (1 + np.exp(-x))


def create_activation(act, sim_kernel=None, use_tf=False):
    """Host (numpy) activation; the device twins live in csrc/sgnn_common.cuh."""
    table = {'relu': relu_np, 'identity': identity_np, 'sigmoid': sigmoid_np, 'tanh': np.tanh}
    if act in table:
        return table[act]
    if act == 'sim_kernel':
        return sim_kernel.dist_to_sim_np
    raise RuntimeError('Unknown activation function {}'.format(act))
