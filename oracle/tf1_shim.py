"""TEST INFRASTRUCTURE -- a NumPy restatement of the TensorFlow 1.x API subset the reference's decode path calls.

The reference's arithmetic lives in TensorFlow 1.11 (README.md:87 of the reference), a third-party dependency that is
neither vendored under /root/reference nor installable here.  This module restates the *published semantics* of the
ops the reference's own graph-construction code uses (model.py, attention_decoder.py, beam_search.py, decode.py,
batcher.py), as a lazily evaluated graph with `Session.run(fetches, feed_dict)`, so that THE REFERENCE'S OWN SOURCE
FILES can be executed unmodified (tests/golden/make_reference_golden.py loads them from /root/reference, in this
container only) and their outputs committed as golden vectors.  What that pins: which ops the reference applies, in
which order, to which operands and with which variable names -- everything the oracle (pgen_oracle.py) restates by
hand.  What it does not pin: TensorFlow's own kernels (Eigen's summation order etc.); each op below follows the TF
1.11 documentation / source behaviour named in its comment, in float32.

Only tests/ and tests/golden/*.py may import this module.  Nothing here is on the product path.

TF 1.11 behaviours restated (file names are TensorFlow's, tensorflow/python/...):
  ops/variable_scope.py   variable_scope (names nest with "/", reuse is inherited, re-entering a captured scope object
                          keeps its full name), get_variable (ValueError on an unexpected duplicate / missing variable)
  ops/rnn_cell_impl.py    LSTMCell: kernel [input + units, 4 units], gates i, j, f, o, forget_bias 1.0, variables
                          "lstm_cell/kernel" and "lstm_cell/bias"; LSTMStateTuple(c, h)
  ops/rnn.py              bidirectional_dynamic_rnn: scopes "bidirectional_rnn/fw", ".../bw"; the backward direction
                          runs on reverse_sequence(inputs, sequence_length) and its outputs are reversed back;
                          dynamic_rnn with sequence_length: zero output and state copied through past the end
  ops/nn_ops.py           conv2d (1 x 1 filters only), softmax (last axis), top_k (sorted, lower index first on ties),
                          xw_plus_b, relu
  ops/array_ops.py        scatter_nd (duplicate indices accumulate, in index order), concat, stack, unstack, tile, ...
  client/session.py       feed_dict may feed ANY tensor (not only placeholders) and nested tuples; values are cast to
                          the fed tensor's dtype
"""
from __future__ import annotations

import collections
import contextlib
import sys
import types

import numpy as np

float32 = np.float32
int32 = np.int32
string = np.object_


# ------------------------------------------------------------------------------------------------
# static shapes
# ------------------------------------------------------------------------------------------------
class Dimension(object):
    def __init__(self, value):
        self.value = None if value is None else int(value)

    def __int__(self):
        return int(self.value)

    __index__ = __int__

    def __eq__(self, other):
        return self.value == (other.value if isinstance(other, Dimension) else other)

    def __hash__(self):
        return hash(self.value)

    def __repr__(self):
        return "Dimension(%r)" % (self.value,)


def _dim(v):
    if isinstance(v, Dimension):
        return v.value
    return None if v is None else int(v)


class TensorShape(object):
    def __init__(self, dims):
        self.dims = None if dims is None else [Dimension(_dim(d)) for d in dims]

    @property
    def ndims(self):
        return None if self.dims is None else len(self.dims)

    def __len__(self):
        return len(self.dims)

    def __getitem__(self, i):
        return self.dims[i]

    def __iter__(self):
        return iter(self.dims)

    def as_list(self):
        return [d.value for d in self.dims]

    def with_rank(self, rank):
        if self.dims is None or len(self.dims) != rank:
            raise ValueError("Shape %r must have rank %d" % (self, rank))
        return self

    def __repr__(self):
        return "TensorShape(%r)" % (None if self.dims is None else self.as_list(),)


def _bcast(a, b):
    """static broadcast of two shape lists (None = unknown dimension)"""
    if a is None or b is None:
        return None
    n = max(len(a), len(b))
    a = [1] * (n - len(a)) + list(a)
    b = [1] * (n - len(b)) + list(b)
    out = []
    for x, y in zip(a, b):
        if x == 1:
            out.append(y)
        elif y == 1:
            out.append(x)
        elif x is None:
            out.append(y)
        else:
            out.append(x)
    return out


# ------------------------------------------------------------------------------------------------
# graph nodes
# ------------------------------------------------------------------------------------------------
_uid = [0]


class Tensor(object):
    __array_priority__ = 1000
    __array_ufunc__ = None            # numpy scalars / arrays defer to the reflected operators below

    def __init__(self, fn, inputs, shape, dtype, name=None):
        self.fn = fn
        self.inputs = list(inputs)
        self._shape = None if shape is None else [_dim(d) for d in shape]
        self.dtype = dtype
        _uid[0] += 1
        self.name = "%s:0" % (name or "op_%d" % _uid[0])

    def get_shape(self):
        return TensorShape(self._shape)

    @property
    def shape(self):
        return self.get_shape()

    def set_shape(self, shape):
        shape = [_dim(d) for d in shape]
        if self._shape is None:
            self._shape = shape
        else:
            assert len(shape) == len(self._shape), (shape, self._shape)
            self._shape = [a if b is None else b for a, b in zip(self._shape, shape)]

    def __iter__(self):
        raise TypeError("Tensor objects are not iterable")

    def __bool__(self):
        raise TypeError("Using a Tensor as a Python bool is not allowed")

    def __repr__(self):
        return "<shim Tensor %s shape=%r>" % (self.name, self._shape)

    # arithmetic keeps float32 (both operands are float32 arrays, or a Python scalar next to one: NEP 50)
    def __add__(self, o): return _binary(np.add, self, o)
    def __radd__(self, o): return _binary(np.add, o, self)
    def __sub__(self, o): return _binary(np.subtract, self, o)
    def __rsub__(self, o): return _binary(np.subtract, o, self)
    def __mul__(self, o): return _binary(np.multiply, self, o)
    def __rmul__(self, o): return _binary(np.multiply, o, self)
    def __truediv__(self, o): return _binary(np.divide, self, o)
    def __rtruediv__(self, o): return _binary(np.divide, o, self)
    __div__ = __truediv__
    __rdiv__ = __rtruediv__

    def __neg__(self):
        return Tensor(lambda a: -a, [self], self._shape, self.dtype)

    def __getitem__(self, idx):
        sh = None
        if self._shape is not None and isinstance(idx, int):
            sh = self._shape[1:]
        return Tensor(lambda a: a[idx], [self], sh, self.dtype)


class Placeholder(Tensor):
    def __init__(self, dtype, shape, name):
        Tensor.__init__(self, None, [], shape, dtype, name)


class Variable(Tensor):
    def __init__(self, initial_value=None, name=None, trainable=True, dtype=None, _full_name=None):
        value = np.asarray(initial_value)
        if dtype is not None:
            value = value.astype(dtype)
        elif value.dtype == np.float64:
            value = value.astype(np.float32)
        elif value.dtype == np.int64:
            value = value.astype(np.int32)
        self.value = value
        full = _full_name
        if full is None:
            scope = get_variable_scope().name
            full = (scope + "/" if scope else "") + (name or "Variable")
            _store()[full] = self
        Tensor.__init__(self, None, [], list(value.shape), value.dtype.type, full)
        self.var_name = full
        self.trainable = trainable

    def assign_value(self, value):
        value = np.asarray(value, dtype=self.value.dtype)
        if value.shape != self.value.shape:
            raise ValueError("variable %s has shape %r, the checkpoint holds %r" % (self.var_name, self.value.shape, value.shape))
        self.value = value


def _const(x):
    if isinstance(x, Tensor):
        return x
    if isinstance(x, Dimension):
        x = x.value
    a = np.asarray(x)
    if a.dtype == np.float64:
        a = a.astype(np.float32)
    elif a.dtype == np.int64:
        a = a.astype(np.int32)
    return Tensor(lambda: a, [], list(a.shape), a.dtype.type)


convert_to_tensor = _const


def _binary(op, a, b):
    # a Python scalar next to a float32 tensor is a float32 constant (TF converts it to the tensor's dtype)
    if isinstance(a, Tensor) and isinstance(b, (int, float)) and a.dtype == np.float32:
        b = np.float32(b)
    if isinstance(b, Tensor) and isinstance(a, (int, float)) and b.dtype == np.float32:
        a = np.float32(a)
    a, b = _const(a), _const(b)
    if a.dtype != b.dtype:
        raise TypeError("tf1_shim: operands of %s have dtypes %r and %r (TensorFlow would refuse them too)" % (op.__name__, a.dtype, b.dtype))
    return Tensor(lambda x, y: op(x, y), [a, b], _bcast(a._shape, b._shape), a.dtype)


def _unary(f, x, dtype=None):
    x = _const(x)
    return Tensor(f, [x], x._shape, dtype or x.dtype)


class _Multi(Tensor):
    """a node with several outputs; `out(i)` selects one"""

    def out(self, i, shape, dtype):
        return Tensor(lambda tup: tup[i], [self], shape, dtype)


# ------------------------------------------------------------------------------------------------
# variables and scopes (ops/variable_scope.py)
# ------------------------------------------------------------------------------------------------
class VariableScope(object):
    def __init__(self, name, reuse):
        self.name = name
        self.reuse = reuse

    def reuse_variables(self):
        self.reuse = True


class _Graph(object):
    def __init__(self):
        self.variables = collections.OrderedDict()
        self.scopes = [VariableScope("", False)]
        self.rng = np.random.RandomState(111)


_graph = [_Graph()]


def reset_default_graph():
    _graph[0] = _Graph()


def _store():
    return _graph[0].variables


def get_variable_scope():
    return _graph[0].scopes[-1]


@contextlib.contextmanager
def variable_scope(name_or_scope, default_name=None, reuse=None, **_unused):
    parent = get_variable_scope()
    if isinstance(name_or_scope, VariableScope):
        # re-entering a captured scope: its full name, reuse inherited from it unless given
        new = VariableScope(name_or_scope.name, name_or_scope.reuse if reuse is None else reuse)
    else:
        name = name_or_scope if name_or_scope is not None else default_name
        full = (parent.name + "/" if parent.name else "") + name
        new = VariableScope(full, parent.reuse if reuse is None else reuse)
    _graph[0].scopes.append(new)
    try:
        yield new
    finally:
        _graph[0].scopes.pop()


def random_uniform_initializer(minval=0, maxval=None, seed=None, dtype=float32):
    return lambda shape: _graph[0].rng.uniform(minval, maxval, size=shape).astype(np.float32)


def truncated_normal_initializer(mean=0.0, stddev=1.0, seed=None, dtype=float32):
    def init(shape):
        x = _graph[0].rng.standard_normal(size=shape)
        bad = np.abs(x) > 2.0
        while bad.any():
            x[bad] = _graph[0].rng.standard_normal(size=int(bad.sum()))
            bad = np.abs(x) > 2.0
        return (mean + stddev * x).astype(np.float32)
    return init


def constant_initializer(value=0.0, dtype=float32):
    return lambda shape: np.full(shape, value, np.float32)


def glorot_uniform_initializer(seed=None, dtype=float32):
    def init(shape):
        if len(shape) == 1:
            fan_in = fan_out = shape[0]
        elif len(shape) == 2:
            fan_in, fan_out = shape
        else:
            rf = int(np.prod(shape[:-2]))
            fan_in, fan_out = shape[-2] * rf, shape[-1] * rf
        lim = np.sqrt(6.0 / (fan_in + fan_out))
        return _graph[0].rng.uniform(-lim, lim, size=shape).astype(np.float32)
    return init


def get_variable(name, shape=None, dtype=float32, initializer=None, trainable=True, **_unused):
    scope = get_variable_scope()
    full = (scope.name + "/" if scope.name else "") + name
    store = _store()
    if full in store:
        if not scope.reuse:
            raise ValueError("Variable %s already exists, disallowed. Did you mean to set reuse=True in VarScope?" % full)
        return store[full]
    if scope.reuse:
        raise ValueError("Variable %s does not exist, or was not created with tf.get_variable()." % full)
    shape = [int(_dim(d)) for d in shape]
    init = initializer or glorot_uniform_initializer()      # get_variable's default initializer
    v = Variable(init(shape), _full_name=full, trainable=trainable)
    store[full] = v
    return v


def global_variables():
    return list(_store().values())


def trainable_variables():
    return [v for v in _store().values() if v.trainable]


# ------------------------------------------------------------------------------------------------
# ops
# ------------------------------------------------------------------------------------------------
def placeholder(dtype, shape=None, name=None):
    return Placeholder(dtype, shape, name)


def _np_dtype(dtype):
    return np.dtype(dtype).type if dtype is not None else np.float32


def _pack(values):
    """a Python list that may hold scalar tensors (e.g. [batch_size, extended_vsize]) -> (inputs, builder)"""
    tens = [(i, v) for i, v in enumerate(values) if isinstance(v, Tensor)]
    static = [None if isinstance(v, Tensor) else int(_dim(v)) for v in values]

    def build(*dyn):
        out = list(static)
        for (i, _), d in zip(tens, dyn):
            out[i] = int(d)
        return out
    return [t for _, t in tens], build, static


def zeros(shape, dtype=float32, name=None):
    dt = _np_dtype(dtype)
    ins, build, static = _pack(list(shape))
    return Tensor(lambda *dyn: np.zeros(build(*dyn), dt), ins, static, dt)


def zeros_like(x):
    return _unary(np.zeros_like, x)


def range(start, limit=None, delta=1, name=None):          # noqa: A001  (tf.range)
    if limit is None:
        start, limit = 0, start
    n = int(np.ceil((limit - start) / float(delta)))
    return Tensor(lambda: np.arange(start, limit, delta, dtype=np.int32), [], [n], np.int32)


def expand_dims(input, axis=None, name=None, dim=None):    # noqa: A002
    if axis is None:
        axis = dim
    x = _const(input)
    sh = None
    if x._shape is not None:
        sh = list(x._shape)
        sh.insert(axis if axis >= 0 else len(sh) + 1 + axis, 1)
    return Tensor(lambda a: np.expand_dims(a, axis), [x], sh, x.dtype)


def shape(input, name=None):                               # noqa: A002
    x = _const(input)
    return Tensor(lambda a: np.asarray(a.shape, np.int32), [x], [None if x._shape is None else len(x._shape)], np.int32)


def tile(input, multiples, name=None):                     # noqa: A002
    x = _const(input)
    ins, build, static = _pack(list(multiples))
    sh = None
    if x._shape is not None:
        sh = [None if (a is None or m is None) else a * m for a, m in zip(x._shape, static)]
    return Tensor(lambda a, *dyn: np.tile(a, build(*dyn)), [x] + ins, sh, x.dtype)


def stack(values, axis=0, name=None):
    vals = [_const(v) for v in values]
    sh = None
    if vals[0]._shape is not None:
        sh = list(vals[0]._shape)
        sh.insert(axis if axis >= 0 else len(sh) + 1 + axis, len(vals))
    return Tensor(lambda *a: np.stack(a, axis=axis), vals, sh, vals[0].dtype)


def unstack(value, num=None, axis=0, name=None):
    x = _const(value)
    n = num if num is not None else x._shape[axis]
    if n is None:
        raise ValueError("Cannot infer num from shape %r" % (x._shape,))
    sh = [d for i, d in enumerate(x._shape) if i != axis]
    return [Tensor(lambda a, i=i: np.take(a, i, axis=axis), [x], sh, x.dtype) for i in np.arange(n)]


def concat(values, axis, name=None):
    vals = [_const(v) for v in values]
    sh = None
    if all(v._shape is not None for v in vals):
        sh = list(vals[0]._shape)
        ax = axis if axis >= 0 else len(sh) + axis
        parts = [v._shape[ax] for v in vals]
        sh[ax] = None if any(p is None for p in parts) else sum(parts)
        for v in vals[1:]:
            sh = [a if a is not None else b for a, b in zip(sh[:ax], v._shape[:ax])] + [sh[ax]] + \
                 [a if a is not None else b for a, b in zip(sh[ax + 1:], v._shape[ax + 1:])]
    return Tensor(lambda *a: np.concatenate(a, axis=axis), vals, sh, vals[0].dtype)


def reshape(tensor, shape, name=None):                     # noqa: A002
    x = _const(tensor)
    ins, build, static = _pack(list(shape))
    sh = [None if (d is None or d < 0) else d for d in static]
    if x._shape is not None and all(d is not None for d in x._shape) and sum(1 for d in static if d is not None and d < 0) == 1 \
            and all(d is not None for d in static):
        known = int(np.prod([d for d in static if d >= 0])) if any(d >= 0 for d in static) else 1
        total = int(np.prod(x._shape)) if x._shape else 1
        sh = [d if d >= 0 else (total // known if known else None) for d in static]
    return Tensor(lambda a, *dyn: np.reshape(a, build(*dyn)), [x] + ins, sh, x.dtype)


def transpose(a, perm=None, name=None):
    x = _const(a)
    sh = None if x._shape is None else (list(reversed(x._shape)) if perm is None else [x._shape[p] for p in perm])
    return Tensor(lambda v: np.transpose(v, perm), [x], sh, x.dtype)


def matmul(a, b, name=None):
    a, b = _const(a), _const(b)
    sh = None
    if a._shape is not None and b._shape is not None:
        sh = [a._shape[0], b._shape[1]]
    return Tensor(lambda x, y: np.matmul(x, y), [a, b], sh, a.dtype)


def reduce_sum(input_tensor, axis=None, keepdims=False, name=None, reduction_indices=None, keep_dims=None):
    x = _const(input_tensor)
    if axis is None:
        axis = reduction_indices
    if keep_dims is not None:
        keepdims = keep_dims
    ax = None if axis is None else tuple(int(a) for a in (axis if isinstance(axis, (list, tuple)) else [axis]))
    sh = None
    if x._shape is not None:
        if ax is None:
            sh = []
        else:
            norm = [a if a >= 0 else len(x._shape) + a for a in ax]
            sh = [(1 if i in norm else d) for i, d in enumerate(x._shape) if keepdims or i not in norm]
    return Tensor(lambda a: np.sum(a, axis=ax, keepdims=keepdims, dtype=a.dtype), [x], sh, x.dtype)


def reduce_mean(input_tensor, axis=None, name=None):
    x = _const(input_tensor)
    return Tensor(lambda a: np.mean(a, axis=axis, dtype=a.dtype), [x], None, x.dtype)


def tanh(x, name=None): return _unary(np.tanh, x)
def log(x, name=None): return _unary(np.log, x)
def exp(x, name=None): return _unary(np.exp, x)
def minimum(x, y, name=None): return _binary(np.minimum, x, y)


def sigmoid(x, name=None):
    return _unary(lambda a: (1.0 / (1.0 + np.exp(-a))).astype(a.dtype), x)


def relu(features, name=None):
    return _unary(lambda a: np.maximum(a, 0).astype(a.dtype), features)


def softmax(logits, axis=-1, name=None, dim=None):
    """nn_ops.softmax: exp(logits - max) / sum over the last axis"""
    def f(a):
        e = np.exp(a - np.max(a, axis=-1, keepdims=True))
        return (e / np.sum(e, axis=-1, keepdims=True, dtype=a.dtype)).astype(a.dtype)
    return _unary(f, logits)


def embedding_lookup(params, ids, name=None):
    p, i = _const(params), _const(ids)
    sh = None if i._shape is None or p._shape is None else list(i._shape) + list(p._shape[1:])
    return Tensor(lambda a, b: a[b], [p, i], sh, p.dtype)


def xw_plus_b(x, weights, biases, name=None):
    x, w, b = _const(x), _const(weights), _const(biases)
    return Tensor(lambda a, m, c: np.matmul(a, m) + c, [x, w, b], [x._shape[0] if x._shape else None, w._shape[1]], x.dtype)


def conv2d(input, filter, strides, padding, name=None, **_unused):     # noqa: A002
    """nn_ops.conv2d restricted to what the reference calls: NHWC input, 1 x 1 filter, stride 1 -- a matmul over the
    channel axis."""
    x, f = _const(input), _const(filter)
    if list(f._shape[:2]) != [1, 1] or list(strides) != [1, 1, 1, 1]:
        raise NotImplementedError("tf1_shim.conv2d: only 1x1 filters with unit strides")
    sh = None if x._shape is None else list(x._shape[:3]) + [f._shape[3]]
    return Tensor(lambda a, w: np.matmul(a, w[0, 0]), [x, f], sh, x.dtype)


def scatter_nd(indices, updates, shape, name=None):        # noqa: A002
    """array_ops.scatter_nd: zeros(shape) with updates ADDED at indices; duplicates accumulate (index order)."""
    idx, upd = _const(indices), _const(updates)
    ins, build, static = _pack(list(shape))

    def f(i, u, *dyn):
        out = np.zeros(build(*dyn), u.dtype)
        np.add.at(out, tuple(np.moveaxis(i, -1, 0)), u)
        return out
    return Tensor(f, [idx, upd] + ins, static, upd.dtype)


TopKV2 = collections.namedtuple("TopKV2", ("values", "indices"))


def top_k(input, k=1, sorted=True, name=None):             # noqa: A002
    """nn_ops.top_k: the k largest entries of the last axis, descending; equal values keep the lower index first."""
    x = _const(input)

    def f(a):
        order = np.argsort(-a, axis=-1, kind="stable")[..., :k]
        return np.take_along_axis(a, order, axis=-1), order.astype(np.int32)
    node = _Multi(f, [x], None, None)
    sh = None if x._shape is None else list(x._shape[:-1]) + [k]
    return TopKV2(node.out(0, sh, x.dtype), node.out(1, sh, np.int32))


# ------------------------------------------------------------------------------------------------
# LSTM (ops/rnn_cell_impl.py, ops/rnn.py)
# ------------------------------------------------------------------------------------------------
LSTMStateTuple = collections.namedtuple("LSTMStateTuple", ("c", "h"))


def _sigm(a):
    return (1.0 / (1.0 + np.exp(-a))).astype(a.dtype)


def _lstm_step(x, c, h, kernel, bias, forget_bias):
    """LSTMCell.call without peepholes / projection: one matmul over [x, h], gates i, j, f, o"""
    z = np.matmul(np.concatenate([x, h], axis=1), kernel) + bias
    i, j, f, o = np.split(z, 4, axis=1)
    c_new = _sigm(f + np.float32(forget_bias)) * c + _sigm(i) * np.tanh(j)
    h_new = _sigm(o) * np.tanh(c_new)
    return c_new.astype(np.float32), h_new.astype(np.float32)


class LSTMCell(object):
    def __init__(self, num_units, initializer=None, state_is_tuple=True, forget_bias=1.0, **_unused):
        self._num_units = num_units
        self._initializer = initializer
        self._forget_bias = forget_bias
        self._kernel = self._bias = None
        assert state_is_tuple

    @property
    def output_size(self):
        return self._num_units

    @property
    def state_size(self):
        return LSTMStateTuple(self._num_units, self._num_units)

    def _build(self, input_depth):
        if self._kernel is None:                    # Layer.__call__ opens the layer's own scope, "lstm_cell"
            with variable_scope("lstm_cell"):
                self._kernel = get_variable("kernel", [input_depth + self._num_units, 4 * self._num_units], initializer=self._initializer)
                self._bias = get_variable("bias", [4 * self._num_units], initializer=constant_initializer(0.0))

    def __call__(self, inputs, state, scope=None):
        x = _const(inputs)
        self._build(x._shape[1])
        fb = self._forget_bias
        node = _Multi(lambda a, c, h, k, b: _lstm_step(a, c, h, k, b, fb), [x, _const(state.c), _const(state.h), self._kernel, self._bias], None, None)
        sh = [x._shape[0], self._num_units]
        c, h = node.out(0, sh, np.float32), node.out(1, sh, np.float32)
        return h, LSTMStateTuple(c, h)


def _reverse_sequence(x, lens):
    out = x.copy()
    for b in np.arange(x.shape[0]):
        n = int(lens[b])
        out[b, :n] = x[b, :n][::-1]
    return out


def _dynamic_rnn(x, lens, kernel, bias, units, forget_bias):
    B, T = x.shape[0], x.shape[1]
    c = np.zeros((B, units), np.float32)
    h = np.zeros((B, units), np.float32)
    outs = np.zeros((B, T, units), np.float32)
    for t in np.arange(T):
        c_new, h_new = _lstm_step(x[:, t, :], c, h, kernel, bias, forget_bias)
        live = (t < lens)[:, None]                  # rnn._rnn_step: zero output and state copied through past the end
        outs[:, t, :] = np.where(live, h_new, np.float32(0))
        c = np.where(live, c_new, c)
        h = np.where(live, h_new, h)
    return outs, c, h


def bidirectional_dynamic_rnn(cell_fw, cell_bw, inputs, sequence_length=None, dtype=None, swap_memory=False, scope=None, **_unused):
    x, lens = _const(inputs), _const(sequence_length)
    depth, units = x._shape[2], cell_fw._num_units
    with variable_scope(scope or "bidirectional_rnn"):
        with variable_scope("fw"):
            cell_fw._build(depth)
        with variable_scope("bw"):
            cell_bw._build(depth)
    fbf, fbb = cell_fw._forget_bias, cell_bw._forget_bias
    fw = _Multi(lambda a, n, k, b: _dynamic_rnn(a, n, k, b, units, fbf), [x, lens, cell_fw._kernel, cell_fw._bias], None, None)

    def run_bw(a, n, k, b):
        outs, c, h = _dynamic_rnn(_reverse_sequence(a, n), n, k, b, units, fbb)
        return _reverse_sequence(outs, n), c, h
    bw = _Multi(run_bw, [x, lens, cell_bw._kernel, cell_bw._bias], None, None)
    osh, ssh = [x._shape[0], x._shape[1], units], [x._shape[0], units]
    outputs = (fw.out(0, osh, np.float32), bw.out(0, osh, np.float32))
    states = (LSTMStateTuple(fw.out(1, ssh, np.float32), fw.out(2, ssh, np.float32)),
              LSTMStateTuple(bw.out(1, ssh, np.float32), bw.out(2, ssh, np.float32)))
    return outputs, states


# ------------------------------------------------------------------------------------------------
# session (client/session.py)
# ------------------------------------------------------------------------------------------------
class OutOfRangeError(Exception):
    pass


class Session(object):
    def __init__(self, target="", graph=None, config=None):
        self.runs = 0

    def _eval(self, t, cache):
        k = id(t)
        if k in cache:
            return cache[k]
        if isinstance(t, Variable):
            out = t.value
        elif isinstance(t, Placeholder):
            raise ValueError("You must feed a value for placeholder tensor %s" % t.name)
        else:
            out = t.fn(*[self._eval(i, cache) for i in t.inputs])
        cache[k] = out
        return out

    def _feed(self, key, value, cache):
        if isinstance(key, tuple):                  # nested structures are fed leaf by leaf
            assert len(key) == len(value)
            for k, v in zip(key, value):
                self._feed(k, v, cache)
            return
        arr = np.asarray(value, dtype=np.dtype(key.dtype) if key.dtype is not None else None)
        if key._shape is not None:
            assert arr.ndim == len(key._shape), "cannot feed shape %r into %r" % (arr.shape, key)
            for d, s in zip(arr.shape, key._shape):
                assert s is None or d == s, "cannot feed shape %r into %r" % (arr.shape, key)
        cache[id(key)] = arr

    def _fetch(self, f, cache):
        if isinstance(f, Tensor):
            return self._eval(f, cache)
        if isinstance(f, dict):
            return {k: self._fetch(v, cache) for k, v in f.items()}
        if isinstance(f, tuple) and hasattr(f, "_fields"):
            return type(f)(*[self._fetch(v, cache) for v in f])
        if isinstance(f, (list, tuple)):
            return [self._fetch(v, cache) for v in f]
        if f is None:
            return None
        raise TypeError("cannot fetch %r" % (f,))

    def run(self, fetches, feed_dict=None):
        cache = {}
        keep = []                                   # fed keys stay alive while their ids are in the cache
        for k, v in (feed_dict or {}).items():
            keep.append(k)
            self._feed(k, v, cache)
        self.runs += 1
        return self._fetch(fetches, cache)

    def close(self):
        pass


# ------------------------------------------------------------------------------------------------
# the module tree the reference imports
# ------------------------------------------------------------------------------------------------
class _Flags(object):
    """tf.app.flags.FLAGS: a plain attribute bag (the generator script sets the flags the reference defines)"""


@contextlib.contextmanager
def _null_context(*_a, **_k):
    yield


def _noop(*_a, **_k):
    return None


def build_modules():
    """-> dict of module name -> module object, to be placed in sys.modules before the reference is imported"""
    me = sys.modules[__name__]
    mods = {}

    def mod(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        mods[name] = m
        parent, _, leaf = name.rpartition(".")
        if parent:
            setattr(mods[parent], leaf, m)
        return m

    tf = mod("tensorflow")
    for n in ("float32", "int32", "string", "placeholder", "variable_scope", "get_variable", "get_variable_scope",
              "random_uniform_initializer", "truncated_normal_initializer", "constant_initializer", "zeros", "zeros_like",
              "range", "expand_dims", "shape", "tile", "stack", "unstack", "concat", "reshape", "transpose", "matmul",
              "reduce_sum", "reduce_mean", "tanh", "log", "exp", "minimum", "sigmoid", "scatter_nd", "Variable",
              "global_variables", "trainable_variables", "Session", "reset_default_graph", "convert_to_tensor"):
        setattr(tf, n, getattr(me, n))
    tf.device = _null_context
    tf.set_random_seed = _noop
    tf.ConfigProto = lambda **kw: types.SimpleNamespace(gpu_options=types.SimpleNamespace(), **kw)
    mod("tensorflow.app")
    mod("tensorflow.app.flags", FLAGS=_Flags())
    mod("tensorflow.logging", info=_noop, warning=_noop, error=_noop, set_verbosity=_noop, INFO=20)
    mod("tensorflow.summary", merge_all=_noop, scalar=_noop, FileWriter=lambda *a, **k: types.SimpleNamespace())
    mod("tensorflow.errors", OutOfRangeError=OutOfRangeError)
    mod("tensorflow.nn", embedding_lookup=embedding_lookup, bidirectional_dynamic_rnn=bidirectional_dynamic_rnn, relu=relu,
        xw_plus_b=xw_plus_b, softmax=softmax, top_k=top_k, conv2d=conv2d, tanh=tanh, sigmoid=sigmoid)
    mod("tensorflow.train", Saver=lambda *a, **k: types.SimpleNamespace(restore=_noop, save=_noop))
    mod("tensorflow.contrib")
    mod("tensorflow.contrib.rnn", LSTMCell=LSTMCell, LSTMStateTuple=LSTMStateTuple)
    mod("tensorflow.contrib.tensorboard")
    mod("tensorflow.contrib.tensorboard.plugins")
    mod("tensorflow.contrib.tensorboard.plugins.projector")
    mod("tensorflow.python")
    mod("tensorflow.python.ops")
    mod("tensorflow.python.ops.variable_scope", variable_scope=variable_scope, get_variable=get_variable,
        get_variable_scope=get_variable_scope)
    mod("tensorflow.python.ops.array_ops", reshape=reshape, zeros=zeros, concat=concat, expand_dims=expand_dims)
    mod("tensorflow.python.ops.nn_ops", conv2d=conv2d, softmax=softmax)
    mod("tensorflow.python.ops.math_ops", reduce_sum=reduce_sum, tanh=tanh, sigmoid=sigmoid)
    mod("tensorflow.python.framework")
    mod("tensorflow.python.framework.errors_impl", OutOfRangeError=OutOfRangeError)
    mod("tensorflow.python.training")
    mod("tensorflow.python.training.monitored_session", WorkerSessionCreator=object, ChiefSessionCreator=object)
    mod("tensorflow.core")
    mod("tensorflow.core.example")
    mod("tensorflow.core.example.example_pb2")
    return mods
