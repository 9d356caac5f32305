"""Eager stand-in for the TensorFlow-1 ops the reference's gaze graphs call — TEST INFRASTRUCTURE ONLY.

Purpose: execute the reference's OWN graph-building code
(``models/gaze_grcn.py``, ``gaze_grcn77.py``, ``gaze_lstm.py``, ``gaze_rnn.py``
``create_loss_and_summary``, ``model_util.py``), imported unmodified from
``/root/reference``, so that the *wiring* of the oracle (which weight meets which
tensor, gate equations, the W_hi reuse, per-timestep BN, loss normalisation) is
pinned by the reference itself.  TensorFlow cannot be installed here, so each
``tf.*`` call lands on a torch-float64 implementation of the documented TF1
semantics.  "Tensors" are plain ``torch.Tensor`` with the three
shape helpers TF code expects patched on; gradients of the reference graph are
then available through torch autograd (the analogue of ``tf.gradients``).

Used only by ``oracle/gen_golden_model.py`` in the build container.
"""
from __future__ import annotations

import contextlib
import math
import sys
import types

import numpy as np
import torch

from oracle import model_oracle as O

DT = torch.float64

# knobs set by the generator (the reference code itself is not touched)
STATE = {
    "rng": np.random.RandomState(0),
    "stddev_scale": 1.0,      # multiplies truncated_normal's stddev (regime 'S' without editing the reference)
    "bn_randomize": False,    # random gamma/beta so per-timestep BN wiring is visible
    "variables": [],          # (name, tensor) in creation order
    "scope": [],
}


class _Shape(object):
    def __init__(self, shape):
        self._s = [int(x) for x in shape]

    def as_list(self):
        return list(self._s)

    def __len__(self):
        return len(self._s)

    def assert_is_compatible_with(self, other):
        other = list(other)
        assert len(other) == len(self._s), (self._s, other)
        for a, b in zip(self._s, other):
            assert b is None or a == b, (self._s, other)


torch.Tensor.get_shape = lambda self: _Shape(self.shape)
torch.Tensor.set_shape = lambda self, s: None


def _t(x):
    return x if isinstance(x, torch.Tensor) else torch.tensor(np.asarray(x), dtype=DT)


def _register(name, t):
    full = "/".join(STATE["scope"] + [name or "Variable"])
    t = t.detach().clone().requires_grad_(True)
    STATE["variables"].append((full, t))
    return t


def Variable(initial_value, name=None, trainable=True, **kw):
    return _register(name, _t(initial_value))


def truncated_normal(shape, dtype=None, stddev=1.0, **kw):
    return _t(O._trunc_normal(STATE["rng"], tuple(shape), stddev * STATE["stddev_scale"]))


def random_uniform(shape, minval=0.0, maxval=1.0, **kw):
    return _t(STATE["rng"].uniform(minval, maxval, size=tuple(shape)))


class _VarScope(object):
    def reuse_variables(self):
        pass


@contextlib.contextmanager
def variable_scope(name=None, *a, **kw):
    STATE["scope"].append(str(name))
    try:
        yield _VarScope()
    finally:
        STATE["scope"].pop()


def get_variable_scope():
    return _VarScope()


def get_variable(name, shape=None, initializer=None, **kw):
    return _register(name, _t(initializer(tuple(shape))))


def transpose(x, perm=None, name=None):
    return x.permute(*perm)


def reshape(x, shape, name=None):
    return x.reshape(*[int(s) for s in shape])


def _is_axis(a):
    return isinstance(a, (int, np.integer))


def concat(*args, **kw):
    """Both TF>=1.0 concat(values, axis) and TF<1.0 concat(concat_dim, values)."""
    values = kw.get("values")
    axis = kw.get("axis", kw.get("concat_dim"))
    for a in args:
        if _is_axis(a):
            axis = a
        else:
            values = a
    return torch.cat(list(values), dim=int(axis))


def split(*args, **kw):
    """TF<1.0 split(split_dim, num_split, value) as used at gaze_lstm.py:112."""
    if _is_axis(args[0]) and _is_axis(args[1]):
        dim, n, value = args
    else:  # TF>=1.0 split(value, num, axis)
        value, n = args[0], args[1]
        dim = args[2] if len(args) > 2 else kw.get("axis", 0)
    return list(torch.chunk(value, int(n), dim=int(dim)))


def stack(values, axis=0, name=None):
    if all(not isinstance(v, torch.Tensor) for v in values):
        return [int(v) for v in values]  # a shape vector (zero_state)
    return torch.stack(list(values), dim=axis)


pack = stack


def zeros(shape, dtype=None, **kw):
    return torch.zeros(*[int(s) for s in shape], dtype=DT)


def placeholder(dtype, shape=None, name=None):
    return torch.zeros(*[int(s) for s in (shape or [])], dtype=DT)


def add(a, b, name=None):
    return a + b


def div(a, b, name=None):
    return a / b


def reduce_sum(x, axis=None, **kw):
    return x.sum() if axis is None else x.sum(dim=axis)


def sigmoid(x, name=None):
    return torch.sigmoid(x)


def tanh(x, name=None):
    return torch.tanh(x)


# ---- tf.nn ----
def _conv2d(x, filter, strides, padding, name=None, **kw):
    assert list(strides) == [1, 1, 1, 1] and padding == "SAME"
    return O.conv2d_same(x, filter)


def _conv2d_transpose(value, filter, output_shape, strides, padding="SAME", name=None, **kw):
    if padding == "SAME" and int(strides[1]) > 1:
        y = O.conv2d_transpose_same_strided(value, filter, int(strides[1]))
    else:
        y = O.conv2d_transpose(value, filter, int(strides[1]), padding)
    assert list(y.shape) == [int(s) for s in output_shape], (y.shape, output_shape)
    return y


def _xw_plus_b(x, w, b, name=None):
    return x @ w + b


def _dropout(x, keep_prob, **kw):
    assert float(keep_prob) == 1.0, "reference graphs are run at keep_prob=1.0 (SURVEY Q2; cascade: inference)"
    return x


def _softmax(x, name=None):
    return torch.softmax(x, dim=-1)


def _softmax_xent(logits=None, labels=None, name=None, **kw):
    return -(labels * torch.log_softmax(logits, dim=-1)).sum(dim=-1)


def _l2_loss(x, name=None):
    return (x * x).sum() / 2


class _RNNCell(object):
    pass


def _batch_normalization(x, **kw):
    """tf.layers.batch_normalization(x), training=False, fresh (gamma,beta) per call."""
    assert not kw.get("training", False)
    c = x.shape[-1]
    n = sum(1 for nm, _ in STATE["variables"] if nm.endswith("/gamma"))
    scope = "batch_normalization" + ("_%d" % n if n else "")
    if STATE["bn_randomize"]:
        g0 = 1.0 + 0.1 * STATE["rng"].standard_normal(size=c)
        b0 = 0.1 * STATE["rng"].standard_normal(size=c)
    else:
        g0, b0 = np.ones(c), np.zeros(c)
    gamma = _register(scope + "/gamma", _t(g0))
    beta = _register(scope + "/beta", _t(b0))
    return gamma * (x - 0.0) / math.sqrt(1.0 + O.BN_EPS) + beta


def _xavier_initializer_conv2d(uniform=True, **kw):
    assert uniform
    return lambda shape: O._xavier_uniform_conv(STATE["rng"], shape)


def _xavier_initializer(uniform=True, **kw):
    assert uniform
    return lambda shape: O._xavier_uniform_fc(STATE["rng"], int(shape[0]), int(shape[1]))


def _constant_initializer(value=0.0, **kw):
    return lambda shape: np.full(tuple(int(s) for s in shape), float(value))


_FC_VARS = {}


def _fully_connected(x, num_output_units, activation_fn=None, weight_init=None, bias_init=None, name=None, **kw):
    """tf.contrib.layers fully_connected of the pre-1.0 API the cascade calls (gaze_grcn_cascade.py:388-411):
    variables <scope>/<name>/weights [in, out] and <scope>/<name>/bias [out], created on the first call and reused
    afterwards (the reference calls reuse_variables() for i > 0); y = x.W + b, then activation_fn."""
    key = "/".join(STATE["scope"] + [str(name)])
    if key not in _FC_VARS or _FC_VARS[key][2] is not STATE["variables"]:
        w = _register(str(name) + "/weights", _t(weight_init((int(x.shape[1]), int(num_output_units)))))
        b = _register(str(name) + "/bias", _t(bias_init((int(num_output_units),))))
        _FC_VARS[key] = (w, b, STATE["variables"])
    w, b, _ = _FC_VARS[key]
    y = x @ w + b
    return activation_fn(y) if activation_fn is not None else y


def _relu(x, name=None):
    return torch.relu(x)


def maximum(a, b, name=None):
    return torch.maximum(a, b)


def _noop(*a, **k):
    return None


def reduce_min(x, axis=None, keep_dims=False, **kw):
    return x.min() if axis is None else torch.amin(x, dim=tuple(axis) if isinstance(axis, (list, tuple)) else axis, keepdim=keep_dims)


def reduce_max(x, axis=None, keep_dims=False, **kw):
    return x.max() if axis is None else torch.amax(x, dim=tuple(axis) if isinstance(axis, (list, tuple)) else axis, keepdim=keep_dims)


def install():
    """Put the stand-in modules into sys.modules so the reference files import unmodified."""
    tf = types.ModuleType("tensorflow")
    for k, v in dict(
        Variable=Variable, truncated_normal=truncated_normal, random_uniform=random_uniform,
        variable_scope=variable_scope, get_variable_scope=get_variable_scope, get_variable=get_variable,
        transpose=transpose, reshape=reshape, concat=concat, split=split, stack=stack, pack=pack,
        zeros=zeros, placeholder=placeholder, add=add, div=div, reduce_sum=reduce_sum,
        sigmoid=sigmoid, tanh=tanh, float32=DT, maximum=maximum, reduce_min=reduce_min, reduce_max=reduce_max, constant_initializer=_constant_initializer,
        scalar_summary=_noop, image_summary=_noop, merge_summary=_noop, get_collection=_noop,
    ).items():
        setattr(tf, k, v)
    nn = types.ModuleType("tensorflow.nn")
    nn.conv2d = _conv2d
    nn.conv2d_transpose = _conv2d_transpose
    nn.xw_plus_b = _xw_plus_b
    nn.dropout = _dropout
    nn.softmax = _softmax
    nn.softmax_cross_entropy_with_logits = _softmax_xent
    nn.l2_loss = _l2_loss
    nn.sigmoid, nn.tanh = sigmoid, tanh
    nn.relu = _relu
    nn.rnn_cell = types.SimpleNamespace(RNNCell=_RNNCell)
    tf.nn = nn
    tf.layers = types.SimpleNamespace(batch_normalization=_batch_normalization)
    inits = types.SimpleNamespace(xavier_initializer_conv2d=_xavier_initializer_conv2d,
                                  xavier_initializer=_xavier_initializer)
    contrib_layers = types.ModuleType("tensorflow.contrib.layers.python.layers")
    contrib_layers.initializers = inits
    contrib_layers.convolution2d = None
    contrib_layers.fully_connected = _fully_connected
    mods = {"tensorflow": tf, "tensorflow.nn": nn,
            "tensorflow.contrib": types.ModuleType("tensorflow.contrib"),
            "tensorflow.contrib.layers": types.ModuleType("tensorflow.contrib.layers"),
            "tensorflow.contrib.layers.python": types.ModuleType("tensorflow.contrib.layers.python"),
            "tensorflow.contrib.layers.python.layers": contrib_layers}
    # non-TF imports of the reference files that are absent / Python-2 only
    import pickle
    mods["cPickle"] = pickle

    class _E(dict):  # easydict.EasyDict
        __getattr__ = dict.get
        __setattr__ = dict.__setitem__
    ed = types.ModuleType("easydict")
    ed.EasyDict = _E
    mods["easydict"] = ed
    cl = types.ModuleType("colorlog")
    import logging
    cl.ColoredFormatter = lambda *a, **k: logging.Formatter("%(message)s")
    mods["colorlog"] = cl
    mods["crc_input_data_seq"] = types.ModuleType("crc_input_data_seq")  # py2 syntax; data loader, not on the path

    class BaseModelConfig(object):  # models/base.py is py2 syntax (print statement at :416); only the ctor is needed
        def __init__(self, *a, **k):
            pass

    class ModelBase(object):
        def __init__(self, *a, **k):
            pass
    mb = types.ModuleType("models.base")
    mb.ModelBase, mb.BaseModelConfig = ModelBase, BaseModelConfig
    mods["models.base"] = mb
    ss = types.ModuleType("models.saliency_shallownet")

    class SaliencyModel(object):
        # models/saliency_shallownet.py (frame-wise ShallowNet) is not on the path: the cascade builds it
        # (gaze_grcn_cascade.py:240-245) but its map never reaches the prediction (the concat is commented out,
        # :366-374) and only feeds an image summary; a zero map of the right shape stands in.
        @staticmethod
        def create_shallownet(images, scope=None, dropout=False, **kw):
            return torch.zeros(int(images.shape[0]), 49, 49, dtype=DT)
    ss.SaliencyModel = SaliencyModel
    mods["models.saliency_shallownet"] = ss
    from oracle.gen_golden_metrics import import_reference_metrics  # skimage stub + evaluation_metrics
    import_reference_metrics()
    for k, v in mods.items():
        sys.modules[k] = v
    for p in ("/root/reference", "/root/reference/models"):
        if p not in sys.path:
            sys.path.insert(0, p)
    import logging as _l
    _l.getLogger("gaze").setLevel(_l.ERROR)


def reset(seed=0, stddev_scale=1.0, bn_randomize=False):
    STATE["rng"] = np.random.RandomState(seed)
    STATE["stddev_scale"] = stddev_scale
    STATE["bn_randomize"] = bn_randomize
    STATE["variables"] = []
    STATE["scope"] = []
