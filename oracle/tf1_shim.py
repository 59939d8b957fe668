"""TEST INFRASTRUCTURE ONLY -- never imported by the product path.

A minimal stand-in for the TensorFlow-1 graph-mode API surface that the
reference's ``model/Siamese`` code touches, evaluated lazily with torch on the
CPU.  Its one purpose: let the UNMODIFIED reference files
(``/root/reference/model/Siamese/{config,inits,layers,layers_factory,models,
model_mse,data_siamese,...}.py``) be imported and executed in this container
(TensorFlow is absent) so that the broadcast quirks and control
flow of the reference (F5 ``layers.py:305-308``, F6 ``model_mse.py:145-151``,
F7 ``model_mse.py:82-92``) are pinned by running the reference's own code, not
by our reading of it.  ``tests/golden/make_golden.py`` installs this module as
``sys.modules['tensorflow']`` and writes the golden vectors.

What is NOT pinned by this: the arithmetic inside each TF op (``tf.matmul``,
``tf.sparse_tensor_dense_matmul``, ``tf.nn.l2_loss``, ``tf.train.AdamOptimizer``
...).  Those live in TensorFlow itself (un-vendored; image
``nvcr.io/nvidia/tensorflow:18.05-py3``, reference ``docker/Dockerfile:1``); they
are restated here from their published definitions.

Only the ops used by the reference are implemented; anything else raises
AttributeError so that silent divergence is impossible.
"""
import contextlib
import math
import os
import sys
import types

import numpy as np
import torch

_DTYPE = torch.float64 if os.environ.get('TF1_SHIM_DTYPE', 'float32') == 'float64' \
    else torch.float32

float32 = 'float32'
int32 = 'int32'
bool = 'bool'  # noqa: A001  (mirrors tf.bool)

_VARIABLES = []
_SCOPE = []


class _Unknown(object):
    """Static shape that is not known to be scalar (compares != ())."""

    def __eq__(self, other):
        return False

    def __ne__(self, other):
        return True


def _conv(x):
    """convert_to_tensor: python numbers / lists of nodes / numpy -> Node."""
    if isinstance(x, Node):
        return x
    if isinstance(x, (list, tuple)) and any(isinstance(e, Node) for e in x):
        return stack(list(x))
    return Const(x)


class Node(object):
    static_scalar = False

    def __init__(self, fn, inputs, static_scalar=False, name=None):
        self.fn = fn
        self.inputs = [_conv(i) for i in inputs]
        self.static_scalar = static_scalar
        self.name = name

    def eval(self, ctx):
        k = id(self)
        if k not in ctx:
            ctx[k] = self.fn(*[i.eval(ctx) for i in self.inputs])
        return ctx[k]

    def get_shape(self):
        return () if self.static_scalar else _Unknown()

    # operators (numpy/TF broadcasting == torch broadcasting)
    def __add__(self, o):
        return Node(lambda a, b: a + b, [self, o])

    def __radd__(self, o):
        return Node(lambda a, b: b + a, [self, o])

    def __sub__(self, o):
        return Node(lambda a, b: a - b, [self, o])

    def __rsub__(self, o):
        return Node(lambda a, b: b - a, [self, o])

    def __mul__(self, o):
        return Node(lambda a, b: a * b, [self, o])

    def __rmul__(self, o):
        return Node(lambda a, b: b * a, [self, o])

    def __truediv__(self, o):
        return Node(lambda a, b: a / b, [self, o])

    def __neg__(self):
        return Node(lambda a: -a, [self])

    def __getitem__(self, idx):
        return Node(lambda a: a[idx], [self])

    def __hash__(self):
        return id(self)

    def __eq__(self, o):
        return self is o


class Const(Node):
    def __init__(self, value):
        self.inputs = []
        self.static_scalar = not hasattr(value, '__len__') and \
            not (isinstance(value, (np.ndarray, torch.Tensor)) and value.ndim > 0)
        if isinstance(value, torch.Tensor):
            self.value = value
        else:
            arr = np.asarray(value)
            if arr.dtype.kind == 'f':
                self.value = torch.as_tensor(arr, dtype=_DTYPE)
            else:
                self.value = torch.as_tensor(arr)
        self.name = None

    def eval(self, ctx):
        return self.value


class Placeholder(Node):
    def __init__(self, dtype, shape=None, default=None):
        self.inputs = []
        self.dtype = dtype
        self.default = default
        self.static_scalar = (shape == ())
        self.name = None

    def eval(self, ctx):
        feeds = ctx['__feeds__']
        if self in feeds:
            v = feeds[self]
        elif self.default is not None:
            v = self.default
        else:
            raise RuntimeError('placeholder not fed')
        arr = np.asarray(v)
        if self.dtype == float32:
            return torch.as_tensor(arr, dtype=_DTYPE)
        return torch.as_tensor(arr.astype(np.int64))


class SparseValue(object):
    """Evaluated sparse tensor: COO triple kept in feed order."""

    def __init__(self, coords, values, shape):
        self.coords = coords
        self.values = values
        self.shape = shape

    def to_dense(self):
        d = torch.zeros(tuple(int(s) for s in self.shape), dtype=self.values.dtype)
        if self.coords.numel():
            d.index_put_((self.coords[:, 0], self.coords[:, 1]), self.values,
                         accumulate=True)
        return d


class SparsePlaceholder(Node):
    def __init__(self, dtype):
        self.inputs = []
        self.dtype = dtype
        self.static_scalar = False
        self.name = None

    def eval(self, ctx):
        coords, values, shape = ctx['__feeds__'][self]
        return SparseValue(torch.as_tensor(np.asarray(coords).astype(np.int64)).reshape(-1, 2),
                           torch.as_tensor(np.asarray(values), dtype=_DTYPE),
                           tuple(shape))

    def __mul__(self, o):  # SparseTensor * scalar  (layers.py:338)
        return Node(lambda a, b: SparseValue(a.coords, a.values * b, a.shape), [self, o])


class _SparseNode(Node):
    def __mul__(self, o):
        return _SparseNode(lambda a, b: SparseValue(a.coords, a.values * b, a.shape),
                           [self, o])


class Variable(Node):
    def __init__(self, initial, name=None):
        self.inputs = []
        self.initial = _conv(initial)
        self.static_scalar = False
        self.name = '/'.join(_SCOPE + [name or 'Variable']) + ':0'
        self.value = None
        _VARIABLES.append(self)

    def initialize(self):
        v = self.initial.eval({'__feeds__': {}})
        self.value = v.clone().detach().to(_DTYPE).requires_grad_(True)

    def assign_value(self, arr):
        self.value = torch.as_tensor(np.asarray(arr), dtype=_DTYPE).clone() \
            .reshape(self.value.shape).requires_grad_(True)

    def numpy(self):
        return self.value.detach().cpu().numpy().copy()

    def eval(self, ctx):
        return self.value


# --------------------------------------------------------------------------
# graph construction helpers
# --------------------------------------------------------------------------
def placeholder(dtype, shape=None):
    return Placeholder(dtype, shape)


def placeholder_with_default(default, shape=None):
    return Placeholder(float32, shape, default=default)


def sparse_placeholder(dtype):
    return SparsePlaceholder(dtype)


@contextlib.contextmanager
def variable_scope(name):
    _SCOPE.append(name)
    try:
        yield
    finally:
        _SCOPE.pop()


class GraphKeys(object):
    GLOBAL_VARIABLES = 'variables'


def get_collection(key, scope=None):
    assert key == GraphKeys.GLOBAL_VARIABLES
    return [v for v in _VARIABLES if scope is None or v.name.startswith(scope)]


def global_variables_initializer():
    def _init():
        for v in _VARIABLES:
            if v.value is None:
                v.initialize()
        return torch.zeros(())

    return Node(_init, [])


# --------------------------------------------------------------------------
# ops (published TF-1 definitions)
# --------------------------------------------------------------------------
def random_uniform(shape, minval=0, maxval=1.0, dtype=None):
    def _f(s):
        dims = [int(x) for x in (s.tolist() if isinstance(s, torch.Tensor) else s)]
        return torch.rand(dims, dtype=_DTYPE) * (maxval - minval) + minval

    return Node(_f, [Const(np.asarray(shape, dtype=np.int64)) if not isinstance(shape, Node)
                     else shape])


def zeros(shape, dtype=None):
    return Const(torch.zeros(list(shape), dtype=_DTYPE))


def ones(shape, dtype=None):
    return Const(torch.ones(list(shape), dtype=_DTYPE))


def floor(x):
    return Node(torch.floor, [x])


def cast(x, dtype=None):
    if dtype == bool:
        return Node(lambda a: a != 0, [x])
    if dtype == float32:
        return Node(lambda a: a.to(_DTYPE), [x])
    raise NotImplementedError(dtype)


def sparse_retain(sp, mask):
    def _f(a, m):
        m = m.reshape(-1)
        return SparseValue(a.coords[m], a.values[m], a.shape)

    return _SparseNode(_f, [sp, mask])


def sparse_tensor_dense_matmul(sp, dense):
    # out[i,:] = sum_e [row_e == i] val_e * dense[col_e,:]
    return Node(lambda a, b: a.to_dense() @ b, [sp, dense])


def matmul(a, b):
    return Node(lambda x, y: x @ y, [a, b])


def multiply(a, b):
    return Node(lambda x, y: x * y, [a, b])


def add_n(xs):
    def _f(*ts):
        out = ts[0]
        for t in ts[1:]:
            out = out + t
        return out

    return Node(_f, list(xs))


def identity(x):
    return Node(lambda a: a, [x])


def sigmoid(x):
    return Node(torch.sigmoid, [x])


def tanh(x):
    return Node(torch.tanh, [x])


def exp(x):
    return Node(torch.exp, [x])


def square(x):
    return Node(lambda a: a * a, [x])


def reduce_mean(x, axis=None):
    if axis is None:
        return Node(lambda a: a.mean(), [x], static_scalar=True)
    return Node(lambda a: a.mean(dim=axis), [x])


def reduce_sum(x, axis=None):
    if axis is None:
        return Node(lambda a: a.sum(), [x], static_scalar=True)
    return Node(lambda a: a.sum(dim=axis), [x])


def reshape(x, shape):
    return Node(lambda a: a.reshape(list(shape)), [x])


def transpose(x):
    return Node(lambda a: a.t() if a.dim() == 2 else a.permute(*reversed(range(a.dim()))), [x])


def concat(xs, axis):
    return Node(lambda *ts: torch.cat(ts, dim=axis), list(xs))


def stack(xs, axis=0):
    return Node(lambda *ts: torch.stack([t if isinstance(t, torch.Tensor) else torch.as_tensor(t, dtype=_DTYPE)
                                         for t in ts], dim=axis), list(xs))


def squeeze(x):
    return Node(lambda a: a.squeeze(), [x])


def shape(x):
    return Node(lambda a: torch.as_tensor(list(a.shape), dtype=torch.int64), [x])


def pad(x, paddings, mode='CONSTANT', constant_values=0):
    assert mode == 'CONSTANT'
    flat = []
    for lo_hi in paddings:
        flat.extend(lo_hi)

    def _f(a, *p):
        p = [int(v) for v in p]
        assert a.dim() == 2 and len(p) == 4
        if p[1] < 0:
            raise ValueError('Paddings must be non-negative')  # tf.pad errors too
        # torch pad order: last dim first
        return torch.nn.functional.pad(a, (p[2], p[3], p[0], p[1]), value=float(constant_values))

    return Node(_f, [x] + flat)


class _NN(object):
    @staticmethod
    def relu(x):
        return Node(torch.relu, [x])

    @staticmethod
    def l2_loss(x):
        return Node(lambda a: (a * a).sum() / 2, [x], static_scalar=True)

    @staticmethod
    def dropout(x, keep_prob):
        # TF-1 nn.dropout: keep_prob == 1 returns x unchanged; otherwise
        # x / keep_prob * floor(keep_prob + U[0,1))
        if isinstance(keep_prob, (int, float)) and keep_prob == 1:
            return _conv(x)
        return Node(lambda a: a / keep_prob * torch.floor(keep_prob + torch.rand_like(a)), [x])


nn = _NN()


class _Summary(object):
    @staticmethod
    def histogram(*a, **k):
        return None

    @staticmethod
    def scalar(*a, **k):
        return None

    @staticmethod
    def merge_all(*a, **k):
        return None

    class FileWriter(object):
        def __init__(self, *a, **k):
            pass

        def add_summary(self, *a, **k):
            pass


summary = _Summary()


class _AdamOptimizer(object):
    """tf.train.AdamOptimizer (TF-1): lr_t = lr*sqrt(1-b2^t)/(1-b1^t);
    m = b1 m + (1-b1) g; v = b2 v + (1-b2) g^2; var -= lr_t * m / (sqrt(v)+eps)."""

    def __init__(self, learning_rate=0.001, beta1=0.9, beta2=0.999, epsilon=1e-8):
        self.lr, self.b1, self.b2, self.eps = learning_rate, beta1, beta2, epsilon
        self.t = 0
        self.m, self.v = {}, {}
        self.last_grads = {}

    def minimize(self, loss):
        opt = self

        def _step(loss_val):
            variables = list(_VARIABLES)
            grads = torch.autograd.grad(loss_val, [v.value for v in variables],
                                        allow_unused=True)
            opt.t += 1
            lr_t = opt.lr * math.sqrt(1 - opt.b2 ** opt.t) / (1 - opt.b1 ** opt.t)
            for var, g in zip(variables, grads):
                if g is None:
                    g = torch.zeros_like(var.value)
                opt.last_grads[var.name] = g.detach().clone()
                k = id(var)
                if k not in opt.m:
                    opt.m[k] = torch.zeros_like(var.value)
                    opt.v[k] = torch.zeros_like(var.value)
                opt.m[k] = opt.b1 * opt.m[k] + (1 - opt.b1) * g
                opt.v[k] = opt.b2 * opt.v[k] + (1 - opt.b2) * g * g
                new = var.value.detach() - lr_t * opt.m[k] / (torch.sqrt(opt.v[k]) + opt.eps)
                var.value = new.clone().requires_grad_(True)
            return torch.zeros(())

        return Node(_step, [loss])


class _Train(object):
    AdamOptimizer = _AdamOptimizer

    class Saver(object):
        def __init__(self, *a, **k):
            pass


train = _Train()


class Session(object):
    def __init__(self, *a, **k):
        self.graph = None

    def run(self, fetches, feed_dict=None):
        ctx = {'__feeds__': feed_dict or {}}
        single = not isinstance(fetches, (list, tuple))
        fl = [fetches] if single else list(fetches)
        # evaluate non-optimizer fetches first so that a fetched loss is the
        # pre-update value (what TF returns when opt_op and loss share a run)
        order = sorted(range(len(fl)), key=lambda i: isinstance(fl[i], Node) and
                       getattr(fl[i].fn, '__name__', '') == '_step')
        outs = [None] * len(fl)
        for i in order:
            f = fl[i]
            if f is None:
                continue
            if isinstance(f, (list, tuple)):
                outs[i] = [self._np(e.eval(ctx)) for e in f]
            else:
                outs[i] = self._np(f.eval(ctx))
        return outs[0] if single else outs

    @staticmethod
    def _np(v):
        if isinstance(v, torch.Tensor):
            v = v.detach().cpu().numpy()
            return v if v.ndim else v[()]
        return v


# --------------------------------------------------------------------------
# tf.app.flags -> absl.flags (absl is what TF-1's tf.app.flags wraps)
# --------------------------------------------------------------------------
class _App(object):
    from absl import flags as _flags
    flags = _flags


app = _App()


def install():
    """Make ``import tensorflow as tf`` resolve to this module."""
    sys.modules['tensorflow'] = sys.modules[__name__]
    return sys.modules[__name__]


def reset():
    del _VARIABLES[:]
    del _SCOPE[:]
