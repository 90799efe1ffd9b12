"""ORACLE -- test infrastructure only (never imported by the product path).

A self-contained NumPy restatement of the reference's training hot path:
`minpy.numpy` primitive -> `Primitive.call` -> tape backward -> `nn.layers` ->
`nn.optim` (dmlc/minpy @2e44927).  Every function cites the reference file:line it
follows (paths relative to /root/reference).  It exists so that the CUDA path can be
checked on a GPU box where /root/reference does not exist.

Pinning: `oracle/gen_golden.py` runs the UNMODIFIED reference (NumPy variant,
`only_numpy` policy) in this container and stores its outputs under `tests/golden/`;
`tests/test_oracle_golden.py` checks this restatement against those vectors and against
the known-answer literals of the reference's own tests (tests/unittest/test_customop.py:48-77,
tests/unittest/test_autograd.py:89-160).  Convolution / pooling have no reference CPU
code (MXNet is not vendored): those two functions are "parity unpinned" and are checked
by central differences only (see conv2d / maxpool2d below).
"""
from __future__ import division

import collections
import itertools
import operator

import numpy as np


# ======================================================================================
#  Array wrapper + tape  (minpy/array.py:26-238, minpy/tape.py:25-257, minpy/core.py:20-90)
# ======================================================================================
class Var(object):
    """Restates `minpy.array.Array`/`Number`: a value with an id and operator overloads
    that redirect to the primitive namespace (array.py:73-221)."""
    _ids = itertools.count(0)
    __array_priority__ = 100.0

    def __init__(self, value):
        self.value = value
        self.id = next(Var._ids)
        self.marked = False

    @property
    def shape(self):
        return np.shape(self.value)

    @property
    def T(self):
        return transpose(self)

    def reshape(self, *shape):
        if len(shape) == 1 and isinstance(shape[0], (tuple, list)):
            shape = shape[0]
        return reshape(self, tuple(shape))

    def dot(self, other):
        return dot(self, other)

    def asnumpy(self):
        return np.array(self.value)

    def __getitem__(self, index):
        return getitem(self, index)

    def __neg__(self): return negative(self)
    def __add__(self, o): return add(self, o)
    def __radd__(self, o): return add(o, self)
    def __sub__(self, o): return subtract(self, o)
    def __rsub__(self, o): return subtract(o, self)
    def __mul__(self, o): return multiply(self, o)
    def __rmul__(self, o): return multiply(o, self)
    def __truediv__(self, o): return true_divide(self, o)
    def __rtruediv__(self, o): return true_divide(o, self)
    def __pow__(self, o): return power(self, o)
    def __rpow__(self, o): return power(o, self)
    def __eq__(self, o): return Var(np.equal(raw(self), raw(o)))
    def __gt__(self, o): return Var(np.greater(raw(self), raw(o)))
    def __lt__(self, o): return Var(np.less(raw(self), raw(o)))
    __hash__ = object.__hash__


def raw(x):
    return x.value if isinstance(x, Var) else x


GradRecord = collections.namedtuple('GradRecord', ['grad_func', 'result', 'owner'])


class Tape(object):
    """Restates minpy/tape.py:25-257 (records, pruning, BFS with ref-counts)."""
    current = None

    def __init__(self):
        self.grads = {}
        self.refcount = collections.defaultdict(int)
        self.records = collections.defaultdict(list)

    def add(self, grad_func, owner, result):  # tape.py:56-84
        self.refcount[owner.id] += 1
        self.records[result.id].append(GradRecord(grad_func, result, owner))

    def _prune(self, start_ids):  # tape.py:113-156
        touched, queue = set(), collections.deque(start_ids)
        while queue:
            cur = queue.popleft()
            if cur in touched:
                continue
            touched.add(cur)
            for rec in self.records[cur]:
                queue.append(rec.owner.id)
        for rid in [r for r in self.records if r not in touched]:
            for rec in self.records[rid]:
                self.refcount[rec.owner.id] -= 1
            self.records.pop(rid)

    def get_gradient(self, origins, target):  # tape.py:158-257
        self.grads[target.id] = 1.0 if np.ndim(target.value) == 0 and not isinstance(target.value, np.ndarray) \
            else np.ones(np.shape(target.value))                                   # tape.py:86-94
        queue = collections.deque([target.id])
        self._prune([target.id])
        origin_ids = set(o.id for o in origins)
        while queue:
            cur = queue.popleft()
            recs = self.records[cur]
            while recs:
                rec = recs.pop()
                g = rec.grad_func(self.grads[cur])
                oid = rec.owner.id
                if oid in self.grads:
                    self.grads[oid] = self.grads[oid] + g                           # tape.py:96-111
                else:
                    self.grads[oid] = g
                self.refcount[oid] -= 1
                if self.refcount[oid] == 0:
                    queue.append(oid)
            self.records.pop(cur, None)
            if cur not in origin_ids:
                self.grads.pop(cur, None)
        return [self.grads.get(o.id, 0.0) for o in origins]                         # tape.py:249-257


def primitive(forward, vjps):
    """Restates `Primitive.call` (minpy/primitive.py:209-377): run forward on raw values;
    for every marked argument build the VJP closure over raw values and record it."""
    def call(*args, **kwargs):
        vals = tuple(raw(a) for a in args)
        res = Var(forward(*vals, **kwargs))
        tape = Tape.current
        if tape is not None:
            marked = [i for i, a in enumerate(args) if isinstance(a, Var) and a.marked]
            if marked:
                res.marked = True
                for i in marked:
                    if i not in vjps:
                        raise ValueError('No partial derivative defined on argument %d' % i)
                    tape.add(vjps[i](res.value, *vals, **kwargs), args[i], res)
        return res
    return call


def grad_and_loss(func, argnum=0):
    """Restates minpy/core.py:20-64."""
    def wrapped(*args):
        arrays = tuple(a if isinstance(a, Var) else Var(a) for a in args)
        argnums = [argnum] if isinstance(argnum, int) else list(argnum)
        Tape.current = tape = Tape()
        try:
            for i in argnums:
                arrays[i].marked = True
            result = func(*arrays)
        finally:
            Tape.current = None
        if not isinstance(result, Var):
            result = Var(result)
        grads = tape.get_gradient([arrays[i] for i in argnums], result)
        for a in arrays:
            a.marked = False
        return (grads[0] if len(grads) == 1 else grads), result.value
    return wrapped


def grad(func, argnum=0):
    f = grad_and_loss(func, argnum)
    return lambda *a: f(*a)[0]


# ======================================================================================
#  Primitives and VJPs  (minpy/array_variants/numpy/numpy_core.py)
# ======================================================================================
def _unbroadcast(ans, x, gradfun):  # numpy_core.py:119-146
    if isinstance(x, np.ndarray):
        shape = x.shape

        def new_fun(g):
            result = gradfun(g)
            while len(shape) < np.ndim(result):
                result = np.sum(result, axis=0)
            for axis, size in enumerate(shape):
                if size == 1:
                    result = np.sum(result, axis=axis, keepdims=True)
            return result
        return new_fun
    elif isinstance(ans, np.ndarray):
        return lambda g: np.sum(gradfun(g))
    return gradfun


_identity = lambda g: g


def _binary(fwd, gx, gy):
    return primitive(fwd, {0: lambda ans, x, y: _unbroadcast(ans, x, gx(ans, x, y)),
                           1: lambda ans, x, y: _unbroadcast(ans, y, gy(ans, x, y))})


add = _binary(np.add, lambda ans, x, y: _identity, lambda ans, x, y: _identity)               # :244-246
subtract = _binary(np.subtract, lambda ans, x, y: _identity, lambda ans, x, y: operator.neg)  # :247-250
multiply = _binary(np.multiply, lambda ans, x, y: (lambda g: g * y), lambda ans, x, y: (lambda g: x * g))  # :240-243
true_divide = _binary(np.true_divide, lambda ans, x, y: (lambda g: g / y),
                      lambda ans, x, y: (lambda g: -g * x / y ** 2))                           # :251-260
divide = true_divide
power = _binary(np.power, lambda ans, x, y: (lambda g: g * y * ans / x),
                lambda ans, x, y: (lambda g: g * np.log(x) * ans))                             # :277-282
maximum = _binary(np.maximum, lambda ans, x, y: (lambda g: g * (x == ans)),
                  lambda ans, x, y: (lambda g: g * (y == ans)))                                # :283-287
minimum = _binary(np.minimum, lambda ans, x, y: (lambda g: g * (x == ans)),
                  lambda ans, x, y: (lambda g: g * (y == ans)))                                # :288-291

dot = primitive(np.dot, {0: lambda ans, a, b: (lambda g: np.dot(g, b.T)),                      # :224-225
                         1: lambda ans, a, b: (lambda g: np.dot(a.T, g))})
tanh = primitive(np.tanh, {0: lambda ans, x: (lambda g: g / np.cosh(x) ** 2)})                 # :228
log = primitive(np.log, {0: lambda ans, x: (lambda g: g / x)})                                 # :229
exp = primitive(np.exp, {0: lambda ans, x: (lambda g: ans * g)})                               # :230
negative = primitive(np.negative, {0: lambda ans, x: operator.neg})                            # :267
transpose = primitive(np.transpose, {0: lambda ans, x: np.transpose})                          # :268
abs = primitive(np.abs, {0: lambda ans, x: (lambda g: np.sign(x) * g)})                        # :269
sqrt = primitive(np.sqrt, {0: lambda ans, x: (lambda g: g * 0.5 / np.sqrt(x))})                # :274
sin = primitive(np.sin, {0: lambda ans, x: (lambda g: g * np.cos(x))})                         # :275
cos = primitive(np.cos, {0: lambda ans, x: (lambda g: -g * np.sin(x))})                        # :276
sign = primitive(np.sign, {0: lambda ans, x: (lambda g: 0.0)})                                 # :270
reshape = primitive(np.reshape, {0: lambda ans, x, shape: (lambda g: np.reshape(g, x.shape))})  # :294-295
expand_dims = primitive(np.expand_dims, {0: lambda ans, x, axis: (lambda g: np.reshape(g, x.shape))})  # :301-302


mod = _binary(np.mod, lambda ans, x, y: _identity,
              lambda ans, x, y: (lambda g: -g * np.floor(x / y)))                              # :261-266
_zero_grad = {0: lambda ans, x: (lambda g: 0.0)}
round = primitive(np.round, _zero_grad)                                                        # :271 (rint-like, zero grad)
ceil = primitive(np.ceil, _zero_grad)                                                          # :272
floor = primitive(np.floor, _zero_grad)                                                        # :273


def _sum_grad(ans, x, axis=None, keepdims=False):  # numpy_core.py:150-162
    if axis is None:
        return lambda g: np.full(x.shape, g)
    axes = [axis] if isinstance(axis, int) else list(axis)
    expanded = list(x.shape)
    for a in axes:
        expanded[a] = 1
    xshape = x.shape
    return lambda g: np.zeros(xshape) + np.reshape(g, expanded)


def _match_shape(a, axis, keepdims):  # numpy_core.py:71-116
    """Function repeating a reduction result back to a.shape (int axes may be negative;
    tuple axes are matched literally, as the reference's `i in axis` does)."""
    if not isinstance(a, np.ndarray):
        return _identity
    shape = a.shape
    if axis is None:
        return lambda g: np.full(shape, g)
    if isinstance(axis, int):
        if keepdims:
            return lambda g: np.repeat(g, shape[axis], axis)
        return lambda g: np.repeat(np.expand_dims(g, axis), shape[axis], axis)
    repeats = [shape[i] if i in axis else 1 for i in range(len(shape))]
    expanded = [shape[i] if i not in axis else 1 for i in range(len(shape))]
    if keepdims:
        return lambda g: np.tile(g, repeats)
    return lambda g: np.tile(np.reshape(g, expanded), repeats)


def _chooser_grad(ans, a, axis=None, keepdims=False):  # numpy_core.py:64-68
    rep = _match_shape(a, axis, keepdims)
    locations = a == rep(ans)
    return lambda g: rep(g) * locations


sum = primitive(np.sum, {0: _sum_grad})                                                        # :233
max = primitive(np.max, {0: _chooser_grad})                                                    # :234
min = primitive(np.min, {0: _chooser_grad})                                                    # :236


def _getitem_grad(ans, x, index):  # numpy_core.py:24-28,292-293
    def g_fun(g):
        ret = np.zeros_like(x)
        np.add.at(ret, index, g)
        return ret
    return g_fun


_getitem_prim = primitive(lambda x, index: x[index], {0: _getitem_grad})


def getitem(x, index):
    if isinstance(index, tuple):
        index = tuple(raw(i) for i in index)
    else:
        index = raw(index)
    return _getitem_prim(x, index)


def _append_fwd(arr, values, axis=None):
    return np.append(arr, values, axis=axis)


append = primitive(_append_fwd, {                                                              # :296-300
    0: lambda ans, arr, values, axis=None: (lambda g: np.split(g, [arr.shape[axis]], axis)[0]),
    1: lambda ans, arr, values, axis=None: (lambda g: np.split(g, [arr.shape[axis]], axis)[1])})


def _sigmoid_fwd(x):  # numpy_core.py:32-43
    pos_mask = (x >= 0)
    neg_mask = (x < 0)
    z = np.zeros_like(x)
    z[pos_mask] = np.exp(-x[pos_mask])
    z[neg_mask] = np.exp(x[neg_mask])
    top = np.ones_like(x)
    top[neg_mask] = z[neg_mask]
    return top / (1 + z)


sigmoid_prim = primitive(_sigmoid_fwd, {0: lambda ans, x: (lambda g: g * ans * (1 - ans))})    # :303


def onehot_encode(indices, out):  # numpy_core.py:47-63 (mutates `out`)
    o = raw(out)
    idx = np.asarray(raw(indices)).astype(int)
    o[np.arange(idx.shape[0]), idx] = 1
    return out


def _softmax_output_fwd(x, _y):  # numpy_core.py:164-168
    probs = np.exp(x - np.max(x, axis=1, keepdims=True))
    probs /= np.sum(probs, axis=1, keepdims=True)
    return probs


softmax_output = primitive(_softmax_output_fwd, {                                              # :170-175
    0: lambda ans, x, y: (lambda _g: (ans - y) / x.shape[0])})


def zeros(shape):
    return Var(np.zeros(shape))


def argmax(x, axis=None):
    return np.argmax(raw(x), axis=axis)


def arange(*args, **kwargs):
    return np.arange(*args, **kwargs)


def array(obj):
    return Var(np.array(raw(obj)))


# ======================================================================================
#  nn.layers  (minpy/nn/layers.py)
# ======================================================================================
def affine(x, w, b):  # layers.py:10-28
    return dot(x, w) + b


def relu(x):  # layers.py:31-42
    return maximum(0, x)


def batchnorm(x, gamma, beta, mode='train', eps=1e-5, momentum=0.9, running_mean=None, running_var=None):
    # layers.py:45-118
    N, D = x.shape
    if running_mean is None:
        running_mean = np.zeros(D)
    if running_var is None:
        running_var = np.zeros(D)
    if mode == 'train':
        mean = sum(x, axis=0) / N
        x_mean = x - expand_dims(mean, axis=0)
        sqr_x_mean = x_mean ** 2
        var = sum(sqr_x_mean, axis=0) / N
        sqrt_var = sqrt(var + eps)
        inv_sqrt_var = 1.0 / sqrt_var
        x_hat = x_mean * expand_dims(inv_sqrt_var, axis=0)
        out = gamma * x_hat + beta
        running_mean = momentum * running_mean + (1.0 - momentum) * mean
        running_var = momentum * running_var + (1.0 - momentum) * var
    elif mode == 'test':
        x_hat = (x - running_mean) / sqrt(running_var + eps)
        out = gamma * x_hat + beta
    else:
        raise ValueError('Invalid forward batchnorm mode "%s"' % mode)
    return out, running_mean, running_var


def dropout(x, prob, mode='train', mask=None):
    """layers.py:121-147.  The reference draws `mask = rand(*x.shape) > prob` from the global
    NumPy RNG; the device path uses Philox, so parity tests pass the mask explicitly."""
    if mode == 'train':
        return x * mask      # not rescaled (layers.py:143-144)
    return x * (1 - prob)    # layers.py:146


def softmax_cross_entropy(prob, label):  # layers.py:176-198
    N, C = prob.shape
    lab = raw(label)
    if len(np.shape(lab)) == 1:
        onehot = np.zeros([N, C])
        onehot[np.arange(N), np.asarray(lab).astype(int)] = 1
    else:
        onehot = lab
    return -sum(log(prob) * onehot) / N


def softmax_loss(x, label):  # layers.py:201-221
    N, C = x.shape
    lab = raw(label)
    if len(np.shape(lab)) == 1:
        onehot = np.zeros([N, C])
        onehot[np.arange(N), np.asarray(lab).astype(int)] = 1
    else:
        onehot = lab
    prob = softmax_output(x, onehot)
    return softmax_cross_entropy(prob, onehot)


def l2_loss(x, label):  # layers.py:224-240
    N, C = x.shape
    lab = raw(label)
    if len(np.shape(lab)) == 1:
        onehot = np.zeros([N, C])
        onehot[np.arange(N), np.asarray(lab).astype(int)] = 1
    else:
        onehot = lab
    return sum((x - onehot) ** 2) / N


def svm_loss(x, y):  # layers.py:150-173
    N = x.shape[0]
    yi = np.asarray(raw(y)).astype(int)
    correct = x[np.arange(N), yi]
    margins = maximum(0, x - correct + 1.0)   # (N,C) - (N,): broadcasts only when N == C, as the reference
    return (sum(margins) - sum(margins[np.arange(N), yi])) / N


def sigmoid(x):  # layers.py:243-253
    return 1 / (1 + exp(-x))


def rnn_step(x, prev_h, Wx, Wh, b):  # layers.py:256-275
    return tanh(dot(x, Wx) + dot(prev_h, Wh) + b)


def lstm_step(x, prev_h, prev_c, Wx, Wh, b):  # layers.py:353-380
    H = prev_c.shape[1]
    a = dot(x, Wx) + dot(prev_h, Wh) + b
    i = sigmoid(a[:, 0:H])
    f = sigmoid(a[:, H:2 * H])
    o = sigmoid(a[:, 2 * H:3 * H])
    g = tanh(a[:, 3 * H:4 * H])
    next_c = f * prev_c + i * g
    next_h = o * tanh(next_c)
    return next_h, next_c


def rnn_temporal(x, h0, Wx, Wh, b):  # layers.py:278-305
    N, T, _ = x.shape
    H = h0.shape[1]
    h = zeros([N, 0, H])
    for t in range(T):
        h_step = rnn_step(x[:, t, :], h0 if t == 0 else h[:, t - 1, :], Wx, Wh, b).reshape(N, 1, H)
        h = append(h, h_step, axis=1)
    return h


def gru_step(x, prev_h, Wx, Wh, b, Wxh, Whh, bh):  # layers.py:308-350
    H = prev_h.shape[1]
    a = sigmoid(dot(x, Wx) + dot(prev_h, Wh) + b)
    r = a[:, 0:H]
    z = a[:, H:2 * H]
    h_m = tanh(dot(x, Wxh) + dot(r * prev_h, Whh) + bh)
    return z * prev_h + (1 - z) * h_m


def lstm_temporal(x, h0, Wx, Wh, b):  # layers.py:383-420
    N, T, _ = x.shape
    H = h0.shape[1]
    c = zeros([N, 0, H])
    h = zeros([N, 0, H])
    for t in range(T):
        h_step, c_step = lstm_step(x[:, t, :], h[:, t - 1, :] if t > 0 else h0,
                                   c[:, t - 1, :] if t > 0 else zeros((N, H)), Wx, Wh, b)
        h = append(h, h_step.reshape(N, 1, H), axis=1)
        c = append(c, c_step.reshape(N, 1, H), axis=1)
    return h


def temporal_affine(x, w, b):  # layers.py:423-441
    N, T, D = x.shape
    M = b.shape[0]
    return x.reshape(N * T, D).dot(w).reshape(N, T, M) + b


def temporal_softmax_loss(x, y, mask):  # layers.py:444-480
    N, T, V = x.shape
    x_flat = x.reshape(N * T, V)
    y_flat = np.asarray(raw(y)).reshape(N * T).astype(int)
    mask_flat = np.asarray(raw(mask)).reshape(N * T)
    probs = exp(x_flat - max(x_flat, axis=1, keepdims=True))
    probs = probs / sum(probs, axis=1, keepdims=True)
    return -sum(mask_flat * log(probs[np.arange(N * T), y_flat])) / N


# ======================================================================================
#  nn.optim  (minpy/nn/optim.py:33-141) -- on raw NumPy arrays
# ======================================================================================
def sgd(w, dw, config):
    config.setdefault('learning_rate', 1e-2)
    return w - config['learning_rate'] * dw, config


def sgd_momentum(w, dw, config):
    config.setdefault('learning_rate', 1e-2)
    config.setdefault('momentum', 0.9)
    v = config.get('velocity', np.zeros_like(w))
    v = v * config['momentum'] - dw * config['learning_rate']
    config['velocity'] = v
    return w + v, config


def rmsprop(x, dx, config):
    config.setdefault('learning_rate', 1e-2)
    config.setdefault('decay_rate', 0.99)
    config.setdefault('epsilon', 1e-8)
    config.setdefault('cache', np.zeros_like(x))
    cache = config['cache'] * config['decay_rate'] + dx ** 2 * (1 - config['decay_rate'])
    config['cache'] = cache
    return x - config['learning_rate'] * dx / (np.sqrt(cache) + config['epsilon']), config


def adam(x, dx, config):
    config.setdefault('learning_rate', 1e-3)
    config.setdefault('beta1', 0.9)
    config.setdefault('beta2', 0.999)
    config.setdefault('epsilon', 1e-8)
    config.setdefault('m', np.zeros_like(x))
    config.setdefault('v', np.zeros_like(x))
    config.setdefault('t', 0)
    t = config['t'] + 1
    m = config['beta1'] * config['m'] + (1 - config['beta1']) * dx
    v = config['beta2'] * config['v'] + (1 - config['beta2']) * (dx ** 2)
    m_ = m / (1 - config['beta1'] ** t)
    v_ = v / (1 - config['beta2'] ** t)
    config['m'], config['v'], config['t'] = m, v, t
    return x - config['learning_rate'] * m_ / (np.sqrt(v_) + config['epsilon']), config


# ======================================================================================
#  nn.init  (minpy/nn/init.py:9-103) -- `z` is the standard-normal draw the rule consumes
# ======================================================================================
def xavier(shape, z):  # init.py:27-34: NORMAL draw scaled by sqrt(6 / (shape[0] + prod(shape[1:])))
    fan_out = shape[0]
    fan_in = int(np.prod(shape[1:]))
    return z * np.sqrt(6.0 / (fan_out + fan_in))


def gaussian(shape, z, config):  # init.py:60-80: defaults mu 0, stdvar 0.001
    return z * config.get('stdvar', 0.001) + config.get('mu', 0.0)


def constant(shape, config):  # init.py:37-57
    return np.ones(shape) * config.get('value', 0.0)


# ======================================================================================
#  Convolution / pooling -- MXNet 0.9 operator semantics as used through core.Function
#  (minpy/core.py:148-190; examples/nn/cnn.py:32-40).  NO reference CPU code exists
#  (MXNet is an un-vendored dependency): PARITY UNPINNED against the reference.  Held to an independent
#  implementation instead -- torch.nn.functional.conv2d / max_pool2d + autograd, CPU float64, incl. the
#  relu-at-zero and first-maximum tie choices (tests/test_conv_oracle_torch.py) -- and to central differences.
# ======================================================================================
def conv2d_fwd(x, w, b, stride=1, pad=0):
    """NCHW cross-correlation: y[n,f,i,j] = b[f] + sum x[n,c,i*s+kh-p,j*s+kw-p] * w[f,c,kh,kw]."""
    N, C, H, W = x.shape
    F, _, KH, KW = w.shape
    OH, OW = (H + 2 * pad - KH) // stride + 1, (W + 2 * pad - KW) // stride + 1
    xp = np.pad(x, ((0, 0), (0, 0), (pad, pad), (pad, pad)))
    y = np.zeros((N, F, OH, OW), dtype=np.result_type(x, w))
    for kh in range(KH):
        for kw in range(KW):
            patch = xp[:, :, kh:kh + stride * OH:stride, kw:kw + stride * OW:stride]
            y += np.einsum('nchw,fc->nfhw', patch, w[:, :, kh, kw])
    if b is not None:
        y += b.reshape(1, F, 1, 1)
    return y


def conv2d_bwd(x, w, dy, stride=1, pad=0):
    N, C, H, W = x.shape
    F, _, KH, KW = w.shape
    OH, OW = dy.shape[2], dy.shape[3]
    xp = np.pad(x, ((0, 0), (0, 0), (pad, pad), (pad, pad)))
    dxp = np.zeros_like(xp, dtype=np.result_type(x, w))
    dw = np.zeros_like(w, dtype=np.result_type(x, w))
    for kh in range(KH):
        for kw in range(KW):
            patch = xp[:, :, kh:kh + stride * OH:stride, kw:kw + stride * OW:stride]
            dw[:, :, kh, kw] = np.einsum('nfhw,nchw->fc', dy, patch)
            dxp[:, :, kh:kh + stride * OH:stride, kw:kw + stride * OW:stride] += np.einsum('nfhw,fc->nchw', dy, w[:, :, kh, kw])
    dx = dxp[:, :, pad:pad + H, pad:pad + W]
    db = dy.sum(axis=(0, 2, 3))
    return dx, dw, db


def maxpool2d_fwd(x, k=2, stride=2):
    """'valid' max pooling; gradient goes to the first maximal element of each window."""
    N, C, H, W = x.shape
    OH, OW = (H - k) // stride + 1, (W - k) // stride + 1
    y = np.full((N, C, OH, OW), -np.inf, dtype=x.dtype)
    arg = np.zeros((N, C, OH, OW), dtype=np.int64)
    for kh in range(k):
        for kw in range(k):
            patch = x[:, :, kh:kh + stride * OH:stride, kw:kw + stride * OW:stride]
            better = patch > y
            y = np.where(better, patch, y)
            hh = (np.arange(OH) * stride + kh).reshape(1, 1, OH, 1)
            ww = (np.arange(OW) * stride + kw).reshape(1, 1, 1, OW)
            arg = np.where(better, hh * W + ww, arg)
    return y, arg


def maxpool2d_bwd(dy, arg, xshape):
    N, C, H, W = xshape
    dx = np.zeros((N * C, H * W), dtype=dy.dtype)
    np.add.at(dx, (np.arange(N * C).reshape(-1, 1), arg.reshape(N * C, -1)), dy.reshape(N * C, -1))
    return dx.reshape(xshape)


# Tape-level wrappers so that CNN workloads can be differentiated by this oracle's tape the
# way `core.Function` exposes a symbol graph as one primitive (minpy/core.py:148-190).
def _conv_prim_fwd(x, w, b, stride=1, pad=0):
    return conv2d_fwd(x, w, b, stride, pad)


conv2d = primitive(_conv_prim_fwd, {
    0: lambda ans, x, w, b, stride=1, pad=0: (lambda g: conv2d_bwd(x, w, g, stride, pad)[0]),
    1: lambda ans, x, w, b, stride=1, pad=0: (lambda g: conv2d_bwd(x, w, g, stride, pad)[1]),
    2: lambda ans, x, w, b, stride=1, pad=0: (lambda g: conv2d_bwd(x, w, g, stride, pad)[2])})


def _pool_vjp(ans, x, k=2, stride=2):
    _, arg = maxpool2d_fwd(x, k, stride)
    return lambda g: maxpool2d_bwd(g, arg, x.shape)


maxpool2d = primitive(lambda x, k=2, stride=2: maxpool2d_fwd(x, k, stride)[0], {0: _pool_vjp})
# MXNet Activation('relu') backward is out_grad * (out > 0): zero at x == 0, unlike layers.relu
relu_mx = primitive(lambda x: np.maximum(x, 0), {0: lambda ans, x: (lambda g: g * (ans > 0))})


def flatten(x):
    return x.reshape(x.shape[0], int(np.prod(x.shape[1:])))


# ---- kink decisions handed in from outside (full-size parity tests, tests/test_fullsize_gpu.py) ----
# A TF32 contraction moves a pre-activation by ~1e-3 relative; the few that sit that close to zero (or
# two pooling candidates that close to each other) land on the other side of relu's / max's kink, and ONE
# such flip changes a whole rank-1 contribution to the weight gradient (at batch 256 that is ~1/sqrt(128)
# of a column).  That is a property of the kink, not an error of a kernel.  The full-size tests therefore
# evaluate the SAME piecewise-linear branch the device took: `keep` is the device's (out > 0) mask,
# `arg` the device's arg-max routing; everything else -- every contraction, normalisation, loss and the
# chain rule through them -- is still this oracle's own float64 arithmetic.
relu_decided = primitive(lambda x, keep: x * keep, {0: lambda ans, x, keep: (lambda g: g * keep)})


def _pool_decided_fwd(x, arg):
    N, C = x.shape[0], x.shape[1]
    flat = x.reshape(N * C, -1)
    return flat[np.arange(N * C).reshape(-1, 1), arg.reshape(N * C, -1)].reshape(arg.shape)


maxpool2d_decided = primitive(_pool_decided_fwd, {0: lambda ans, x, arg: (lambda g: maxpool2d_bwd(g, arg, x.shape))})
