"""Primitives and VJPs of the CUDA array variant.

Plugs into the reference's variant interface (`register_primitives(reg, prim_wrapper)`,
`def_grads(prims)`: minpy/array_variants/mxnet/mxnet_core.py:100-170 is the template) with
functions that enqueue kernels of libminpy_cuda.so through `minpy_b200.ops`.

Two things go beyond the template:

* VJPs of elementwise primitives do not launch kernels.  They append a step to a lazy
  `ChainGrad` recipe; the rewritten tape evaluates a whole run of them with one
  `mc_vjp_chain` launch (minpy/tape.py, minpy_b200/lazy.py).
* `fused_*` primitives (affine, relu, batchnorm, dropout, softmax loss, LSTM step,
  optimizer updates) are whole-layer kernels that `minpy.nn.layers` / `minpy.nn.optim`
  pick when they are bound in the namespace; their arithmetic restates the primitive
  compositions of the reference's nn/layers.py and nn/optim.py (cited per function).
"""
from __future__ import absolute_import
from __future__ import division

import numpy as np

from minpy_b200 import ops
from minpy_b200._lib import parse_enums
from minpy_b200.cuda_array import CudaArray
from minpy_b200.lazy import ChainGrad, LazyArray, as_chain, concrete, drop_step, DROPOUT, FULL, ROW, COL, ONE

E = parse_enums()
_prod = ops._prod  # pylint: disable=protected-access


def _is_arr(x):
    return isinstance(x, CudaArray)


def _shape(x):
    return x.shape if _is_arr(x) else np.shape(x)


# ======================================================================================
#  operand addressing and un-broadcasting for chain steps
# ======================================================================================
def _opd(v, shape):
    """How array `v` (broadcastable to `shape`) is addressed inside a chain over `shape`:
    returns (array, mode, inner)."""
    if v.dtype != ops._F32:  # pylint: disable=protected-access
        v = v.astype(ops._F32)  # pylint: disable=protected-access
    if v.shape == shape or v.size == _prod(shape):
        return v, FULL, 1
    if v.size == 1:
        return v, ONE, 1
    vs = (1,) * (len(shape) - v.ndim) + v.shape
    k = len(shape)
    while k > 0 and vs[k - 1] == shape[k - 1]:
        k -= 1
    if all(d == 1 for d in vs[:k]):                       # trailing block matches: row vector
        return v, ROW, _prod(shape[k:])
    k = 0
    while k < len(shape) and vs[k] == shape[k]:
        k += 1
    if all(d == 1 for d in vs[k:]):                       # leading block matches: column vector
        return v, COL, _prod(shape[k:])
    return ops.broadcast_to(v, shape), FULL, 1


def _fold(chain, operand):
    """Gradient recipe in the broadcast shape -> gradient of `operand` (its own shape)."""
    if _is_arr(operand):
        if operand.shape == chain.shape:
            return chain
        return ops.unbroadcast(chain.materialize(), operand.shape)
    return ops.reduce(E['MC_RED_SUM'], chain.materialize())    # scalar operand: sum everything


def _step1(chain, op_arr, op_scalar, v, shape):
    """Append `op_arr` with operand v (array) or `op_scalar` with scalar v."""
    if _is_arr(v):
        a, mode, inner = _opd(v, shape)
        return chain.then(E[op_arr], a, mode, inner=inner)
    return chain.then(E[op_scalar], scalar=float(v))


def _step2(chain, op, a, b, shape, scalar=0.0):
    a, am, ai = _opd(a, shape)
    b, bm, bi = _opd(b, shape)
    if am in (ROW, COL) and bm in (ROW, COL) and ai != bi:
        b, bm, bi = ops.broadcast_to(b, shape), FULL, 1
    inner = ai if am in (ROW, COL) else bi
    return chain.then(E[op], a, am, b, bm, scalar, inner)


def _binary_vjps(name):
    """(d/dx maker, d/dy maker) for a broadcasting binary ufunc."""
    def dx(ans, x, y):
        shape = ans.shape

        def vjp(g):
            c = as_chain(g, shape)
            if name in ('add', 'subtract', 'mod'):
                pass
            elif name == 'multiply':
                c = _step1(c, 'MC_VJP_MUL_A', 'MC_VJP_MUL_S', y, shape)
            elif name in ('divide', 'true_divide'):
                c = _step1(c, 'MC_VJP_DIV_A', 'MC_VJP_DIV_S', y, shape)
            elif name == 'power':
                # g * y * x^(y-1): the MXNet variant's form (mxnet_core.py:160-162), finite at
                # x == 0 where the NumPy variant's g*y*ans/x is NaN
                if _is_arr(y):
                    c = _step2(c, 'MC_VJP_POW_X', x, y, shape)
                else:
                    a, mode, inner = _opd(x, shape)
                    c = c.then(E['MC_VJP_POW_XS'], a, mode, scalar=float(y), inner=inner)
            elif name in ('maximum', 'minimum'):
                c = _step2(c, 'MC_VJP_EQ_AB', x, ans, shape)
            return _fold(c, x)
        return vjp

    def dy(ans, x, y):
        shape = ans.shape

        def vjp(g):
            c = as_chain(g, shape)
            if name == 'add':
                pass
            elif name == 'subtract':
                c = c.then(E['MC_VJP_NEG'])
            elif name == 'multiply':
                c = _step1(c, 'MC_VJP_MUL_A', 'MC_VJP_MUL_S', x, shape)
            elif name in ('divide', 'true_divide'):
                if _is_arr(x):
                    c = _step2(c, 'MC_VJP_DIV_Y', x, y, shape)
                else:
                    a, mode, inner = _opd(y, shape)
                    c = c.then(E['MC_VJP_S_DIV_Y'], a, mode, scalar=float(x), inner=inner)
            elif name == 'mod':
                q = ops.unary(E['MC_FLOOR'], ops.binary(E['MC_DIV'], x, y))
                c = c.then(E['MC_VJP_MUL_A'], *_opd(q, shape)[:2]).then(E['MC_VJP_NEG'])
            elif name == 'power':
                if _is_arr(x):
                    c = _step2(c, 'MC_VJP_POW_Y', x, ans, shape)
                else:
                    c = c.then(E['MC_VJP_POW_SY'], ans, FULL, scalar=float(x))
            elif name in ('maximum', 'minimum'):
                c = _step2(c, 'MC_VJP_EQ_AB', y, ans, shape)
            return _fold(c, y)
        return vjp
    return dx, dy


# unary: name -> (forward opcode, VJP step opcode or None for zero gradient, operand: 'x'|'ans'|None)
_UNARY = {
    'negative': ('MC_NEG', 'MC_VJP_NEG', None),
    'abs': ('MC_ABS', 'MC_VJP_ABS_X', 'x'),
    'absolute': ('MC_ABS', 'MC_VJP_ABS_X', 'x'),
    'sign': ('MC_SIGN', None, None),
    'round': ('MC_ROUND', None, None),
    'ceil': ('MC_CEIL', None, None),
    'floor': ('MC_FLOOR', None, None),
    'sqrt': ('MC_SQRT', 'MC_VJP_SQRT_ANS', 'ans'),
    'exp': ('MC_EXP', 'MC_VJP_MUL_A', 'ans'),
    'log': ('MC_LOG', 'MC_VJP_DIV_A', 'x'),
    'sin': ('MC_SIN', 'MC_VJP_SIN_X', 'x'),
    'cos': ('MC_COS', 'MC_VJP_COS_X', 'x'),
    'tanh': ('MC_TANH', 'MC_VJP_TANH_ANS', 'ans'),
    'sigmoid': ('MC_SIGMOID', 'MC_VJP_SIGMOID_ANS', 'ans'),
    'reciprocal': ('MC_RECIP', 'MC_VJP_RECIP_ANS', 'ans'),
    'cosh': ('MC_COSH', None, None),
    'logical_not': ('MC_LOGICAL_NOT', None, None),
}

_BINARY = {
    'add': 'MC_ADD', 'subtract': 'MC_SUB', 'multiply': 'MC_MUL', 'divide': 'MC_DIV',
    'true_divide': 'MC_DIV', 'power': 'MC_POW', 'maximum': 'MC_MAX', 'minimum': 'MC_MIN', 'mod': 'MC_MOD',
}
_COMPARE = {
    'equal': 'MC_EQ', 'not_equal': 'MC_NE', 'less': 'MC_LT', 'greater': 'MC_GT',
    'less_equal': 'MC_LE', 'greater_equal': 'MC_GE',
}


def _named(fn, name):
    fn.__name__ = name
    return fn


def _make_unary(name, opcode):
    def fwd(x):
        if not _is_arr(x):
            return getattr(np, name if name != 'sigmoid' else 'tanh')(x) if name != 'sigmoid' else 1.0 / (1.0 + np.exp(-x))
        return ops.unary(E[opcode], x)
    return _named(fwd, name)


def _make_binary(name, opcode):
    def fwd(x, y):
        if not _is_arr(x) and not _is_arr(y):
            return getattr(np, name)(x, y)               # scalar op scalar stays a host scalar
        return ops.binary(E[opcode], x, y)
    return _named(fwd, name)


def square(x):
    return ops.unary(E['MC_SQUARE'], x)


# ======================================================================================
#  reductions
# ======================================================================================
def _reduce_fn(name, opcode):
    def fwd(x, axis=None, out=None, keepdims=False):
        if out is not None:
            raise ValueError('Out option is not supported.')
        if not _is_arr(x):
            return getattr(np, name)(x, axis=axis, keepdims=keepdims)
        return ops.reduce(E[opcode], x, axis, keepdims)
    return _named(fwd, name)


def _single_run(shape, axis):
    """(outer, red, inner) when the reduced axes are adjacent, else None."""
    axes = ops._norm_axes(axis, len(shape))  # pylint: disable=protected-access
    if not axes or list(axes) != list(range(axes[0], axes[-1] + 1)):
        return None
    return _prod(shape[:axes[0]]), _prod(shape[axes[0]:axes[-1] + 1]), _prod(shape[axes[-1] + 1:])


def _sum_vjp(ans, x, axis=None, out=None, keepdims=False):  # pylint: disable=unused-argument
    """Broadcast g back over the reduced axes (numpy_core.py:150-162) -- as an addressing mode
    of the chain kernel when the layout allows, so it costs no pass of its own."""
    xshape = x.shape

    def vjp(g):
        if isinstance(g, ChainGrad):
            g = g.materialize()
        if not _is_arr(g):
            return ChainGrad(xshape, None, float(g))
        if g.size == 1:
            return ChainGrad(xshape, g, g_mode=ONE)
        run = _single_run(xshape, axis)
        if run is not None:
            outer, red, inner = run
            if outer == 1:
                return ChainGrad(xshape, g, g_mode=ROW, g_inner=inner)
            if inner == 1:
                return ChainGrad(xshape, g, g_mode=COL, g_inner=red)
        return ops.expand_reduced(g, xshape, axis, keepdims)
    return vjp


def _extremum_vjp(ans, x, axis=None, out=None, keepdims=False):  # pylint: disable=unused-argument
    """g at every position equal to the extremum, ties included (numpy_core.py:66-70)."""
    def vjp(g):
        g = concrete(g, ans.shape)
        return ops.select_grad(x, ans, g, axis, keepdims)
    return vjp


def argmax(x, axis=None, out=None):
    if out is not None:
        raise ValueError('Out option is not supported.')
    return ops.argmax(x, axis)


def mean(x, axis=None, keepdims=False):
    s = ops.reduce(E['MC_RED_SUM'], x, axis, keepdims)
    count = x.size // max(1, s.size)
    return ops.binary(E['MC_MUL'], s, 1.0 / count)


def count_nonzero(x, axis=None):
    nz = ops.binary(E['MC_NE'], x, 0.0)
    return ops.reduce(E['MC_RED_SUM'], nz, axis)


# ======================================================================================
#  shape ops, indexing, creation
# ======================================================================================
def reshape(x, shape, order='C'):
    if order != 'C':
        raise NotImplementedError('Orders other than C are not currently supported.')
    if isinstance(shape, (int, np.integer)):
        shape = (int(shape),)
    return ops.as_cuda(x).reshape(tuple(shape))


def _reshape_back(ans, x, *args, **kwargs):  # pylint: disable=unused-argument
    xshape = x.shape

    def vjp(g):
        if isinstance(g, ChainGrad):
            return g.reshaped(xshape)
        if _is_arr(g):
            return g.reshape(xshape)
        return g
    return vjp


def transpose(x, axes=None):
    return ops.transpose(x, axes)


def _transpose_vjp(ans, x, axes=None):  # pylint: disable=unused-argument
    inverse = None if axes is None else tuple(int(i) for i in np.argsort(axes))
    return lambda g: ops.transpose(concrete(g, ans.shape), inverse)


def expand_dims(x, axis):
    return ops.expand_dims(x, axis)


def squeeze(x, axis=None):
    if axis is None:
        shape = tuple(s for s in x.shape if s != 1)
    else:
        axes = ops._norm_axes(axis, x.ndim)  # pylint: disable=protected-access
        shape = tuple(s for i, s in enumerate(x.shape) if i not in axes)
    return x.reshape(shape)


def ravel(x):
    return x.reshape((x.size,))


def broadcast_to(x, shape):
    return ops.broadcast_to(x, shape)


def append(arr, values, axis=None):
    arr, values = ops.as_f32(arr), ops.as_f32(values)
    if axis is None:
        return ops.concat2(arr.reshape((arr.size,)), values.reshape((values.size,)), 0)
    return ops.concat2(arr, values, axis)


def _append_vjp(pos):
    def maker(ans, arr, values, axis=None):
        def vjp(g):
            g = concrete(g, ans.shape)
            if axis is None:
                n = arr.size
                piece = ops.getitem(g, slice(0, n) if pos == 0 else slice(n, None))
                return piece.reshape((arr, values)[pos].shape)
            ax = axis + ans.ndim if axis < 0 else axis
            n = arr.shape[ax]
            index = (slice(None),) * ax + ((slice(0, n) if pos == 0 else slice(n, None)),)
            return ops.getitem(g, index)
        return vjp
    return maker


def concatenate(arrays, axis=0):
    arrays = [ops.as_f32(a) for a in arrays]
    out = arrays[0]
    for nxt in arrays[1:]:
        out = ops.concat2(out, nxt, axis)
    return out


def stack(arrays, axis=0):
    arrays = [ops.expand_dims(ops.as_f32(a), axis) for a in arrays]
    return concatenate(arrays, axis)


def _minpy_getitem(arr, index):
    return ops.getitem(arr, index)


def _getitem_vjp(ans, x, index):
    return lambda g: ops.getitem_grad(x, index, concrete(g, ans.shape))


def _shape_arg(shape, rest=()):
    if isinstance(shape, (int, np.integer)):
        return (int(shape),) + tuple(int(r) for r in rest)
    return tuple(int(s) for s in shape)


def zeros(shape, dtype=None, order='C'):  # pylint: disable=unused-argument
    return ops.zeros(_shape_arg(shape))


def ones(shape, dtype=None, order='C'):  # pylint: disable=unused-argument
    return ops.full(_shape_arg(shape), 1.0)


def empty(shape, dtype=None, order='C'):  # pylint: disable=unused-argument
    return ops.empty(_shape_arg(shape))


def full(shape, fill_value, dtype=None):  # pylint: disable=unused-argument
    return ops.full(_shape_arg(shape), float(fill_value))


def zeros_like(a, dtype=None):  # pylint: disable=unused-argument
    return ops.zeros(_shape(a))


def ones_like(a, dtype=None):  # pylint: disable=unused-argument
    return ops.full(_shape(a), 1.0)


def array(obj, dtype=None, copy=True):  # pylint: disable=unused-argument,redefined-outer-name
    if _is_arr(obj):
        return obj.copy()
    return CudaArray.from_numpy(np.asarray(obj))


def arange(*args, **kwargs):
    return ops.arange(*args, **kwargs)


def eye(n, m=None, k=0, dtype=None):  # pylint: disable=unused-argument
    return CudaArray.from_numpy(np.eye(n, m, k, dtype=np.float32))


# ======================================================================================
#  dot, softmax_output, onehot_encode
# ======================================================================================
def dot(a, b):
    return ops.dot(a, b)


def _as2d(v, lead_is_row):
    """1-D operands of dot as matrices: a (K,) -> (1,K); b (K,) -> (K,1)."""
    if v.ndim == 1:
        return v.reshape((1, v.shape[0]) if lead_is_row else (v.shape[0], 1))
    if v.ndim > 2:
        return v.reshape((_prod(v.shape[:-1]), v.shape[-1]))
    return v


def _dot_vjp_a(ans, a, b):
    """dA = g . B^T (numpy_core.py:224)."""
    def vjp(g):
        if not _is_arr(b):
            return _fold(as_chain(g, ans.shape).then(E['MC_VJP_MUL_S'], scalar=float(b)), a)
        if not _is_arr(a):
            return ops.reduce(E['MC_RED_SUM'], ops.binary(E['MC_MUL'], concrete(g, ans.shape), b))
        gm = concrete(g, ans.shape)
        a2, b2 = _as2d(a, True), _as2d(b, False)
        g2 = gm.reshape((a2.shape[0], b2.shape[1]))
        return ops.gemm(g2, b2, trans_b=True).reshape(a.shape)
    return vjp


def _dot_vjp_b(ans, a, b):
    """dB = A^T . g (numpy_core.py:225)."""
    def vjp(g):
        if not _is_arr(a):
            return _fold(as_chain(g, ans.shape).then(E['MC_VJP_MUL_S'], scalar=float(a)), b)
        if not _is_arr(b):
            return ops.reduce(E['MC_RED_SUM'], ops.binary(E['MC_MUL'], concrete(g, ans.shape), a))
        gm = concrete(g, ans.shape)
        a2, b2 = _as2d(a, True), _as2d(b, False)
        g2 = gm.reshape((a2.shape[0], b2.shape[1]))
        return ops.gemm(a2, g2, trans_a=True).reshape(b.shape)
    return vjp


def softmax_output(x, _label):
    return ops.softmax_rows(x)


def _softmax_output_vjp(ans, x, y):
    """(p - y) / N; the upstream gradient is ignored, exactly as numpy_core.py:170-175."""
    def vjp(_g):
        out = CudaArray.empty(ans.shape)
        if _is_arr(y) and y.ndim == 1:
            ops.L().mc_softmax_ce_bwd(ans.ptr, ops.as_i32(y).ptr, None, out.ptr, ans.shape[0], ans.shape[1])
        else:
            ops.L().mc_softmax_ce_bwd(ans.ptr, None, ops.as_f32(y).ptr, out.ptr, ans.shape[0], ans.shape[1])
        return out
    return vjp


def onehot_encode(indices, out):
    return ops.onehot_encode(indices, out)


# ======================================================================================
#  fused layers (picked by minpy.nn.layers when bound)
# ======================================================================================
import os as _os

# Forward fusion across layer calls (MINPY_FUSE_FORWARD=0 turns it off): `layers.affine` and
# `layers.batchnorm` return LazyArrays, so that a relu / dropout called on them next is absorbed into the
# producing kernel -- the hidden block affine -> relu -> batchnorm -> dropout of mlp_bn_dropout.py:41-54 is
# TWO launches forward (GEMM with bias + relu epilogue; batchnorm with the dropout mask applied at the store)
# and its elementwise backward ONE (dropout + batchnorm + relu VJPs and the bias gradient, see _batchnorm_vjp).
_FUSE_FORWARD = _os.environ.get('MINPY_FUSE_FORWARD', '1') != '0'
_RELU_SIGNED = ('relu_signed',)     # tag of a relu output whose zeros carry the sign of the pre-activation


def _relu_done(g, relu_out):
    """True when gradient `g` (of `relu_out`) already had relu's mask applied by batchnorm's fused VJP."""
    tag = g._tag if isinstance(g, CudaArray) else None  # pylint: disable=protected-access
    return tag is not None and tag[0] == 'relu_done' and tag[1] is relu_out


def fused_affine(x, w, b):
    """dot(x, w) + b in one GEMM with a bias epilogue (nn/layers.py:27); evaluated lazily so that a
    relu called on the result rides in the same epilogue."""
    x = ops.as_f32(x)
    b1 = b.reshape((b.size,))
    if not _FUSE_FORWARD or x.ndim != 2 or w.ndim != 2 or x.shape[1] != w.shape[0] or b1.size != w.shape[1]:
        return ops.gemm(x, w, bias=b1)
    return LazyArray((x.shape[0], w.shape[1]), lambda self: self.bind(ops.gemm(x, w, bias=b1)), saved=('affine', x, w, b1))


def _affine_vjps():
    # (`ans` may be a LazyArray that was never evaluated: only its shape is used)
    def dx(ans, x, w, b):  # pylint: disable=unused-argument
        def vjp(g):
            tag = g._tag if isinstance(g, CudaArray) else None  # pylint: disable=protected-access
            if tag is not None and tag[0] == 'head_done' and tag[1] is ans:
                return tag[2]                           # computed by the classifier-head kernel (fused_softmax_loss)
            return ops.gemm(concrete(g, ans.shape), w, trans_b=True)
        return vjp

    def dw(ans, x, w, b):  # pylint: disable=unused-argument
        def vjp(g):
            # x^T . g is read only by the update rule: on the side stream, beside the dX chain (ops.side_stream)
            xf, gc = ops.as_f32(x), concrete(g, ans.shape)
            with ops.side_stream(xf, gc):
                out = ops.gemm(xf, gc, trans_a=True)
            ops.side_keep(out)
            return out
        return vjp

    def db(ans, x, w, b):  # pylint: disable=unused-argument
        def vjp(g):
            tag = g._tag if isinstance(g, CudaArray) else None  # pylint: disable=protected-access
            if tag is not None and tag[0] == 'relu_done':
                return tag[2].reshape(b.shape)          # column sums computed by batchnorm's fused VJP
            gc = concrete(g, ans.shape)
            with ops.side_stream(gc):                   # read only by the update rule: beside the dX chain
                out = ops.reduce(E['MC_RED_SUM'], gc, 0)
            ops.side_keep(out)
            return out.reshape(b.shape)
        return vjp
    return dx, dw, db


def fused_relu(x):
    """maximum(0, x) (nn/layers.py:41).  On a pending affine: the GEMM runs now with bias + relu in its
    epilogue (relu mode 2: zeros keep the sign of the pre-activation, which the VJP needs)."""
    if type(x) is LazyArray and x.pending() and type(x._saved) is tuple and x._saved[0] == 'affine':  # pylint: disable=protected-access,unidiomatic-typecheck
        _, xx, w, b1 = x._saved  # pylint: disable=protected-access
        out = ops.gemm(xx, w, bias=b1, relu=2)
        out._tag = _RELU_SIGNED  # pylint: disable=protected-access
        return out
    return ops.unary(E['MC_RELU'], x)


def _relu_vjp(ans, x):
    if ans._tag is _RELU_SIGNED:  # pylint: disable=protected-access
        # fused with its affine: x was never computed; g * !signbit(ans) == g * (x >= 0)
        def vjp(g):
            if _relu_done(g, ans):
                return g
            return as_chain(g, ans.shape).then(E['MC_VJP_SIGNPOS_A'], ans, FULL)
        return vjp
    # g * (x == ans): passes g where x >= 0, including x == 0 (numpy_core.py:284-286)
    return lambda g: as_chain(g, ans.shape).then(E['MC_VJP_GE0_A'], x, FULL)


class _BnPending(object):
    """One train-mode batchnorm call whose kernel has not run yet: its three results (out, new running
    mean, new running variance) are LazyArrays that share this record."""
    __slots__ = ('x', 'gamma', 'beta', 'rm', 'rv', 'eps', 'momentum', 'out', 'new_rm', 'new_rv', 'save_mean',
                 'save_invstd', 'stats_done', 'drop')

    def __init__(self, x, gamma, beta, rm, rv, eps, momentum):
        self.x, self.gamma, self.beta, self.rm, self.rv = x, gamma, beta, rm, rv
        self.eps, self.momentum = float(eps), float(momentum)
        self.save_mean = self.save_invstd = None
        self.stats_done = False
        self.drop = None            # (prob, seed, offset) of the dropout fused behind it

    def _launch(self, out, drop):
        lib = ops.L()
        x = self.x
        N, D = x.shape
        from minpy_b200.cuda_array import _Buffer
        if _Buffer._active_pool is not None and not self.stats_done:  # pylint: disable=protected-access
            # step-graph capture: the running statistics are updated in place (each column is read and written by
            # one thread), so the graph needs no copy-back node per statistic (step_graph.py)
            new_rm, new_rv = self.rm, self.rv
        else:
            new_rm, new_rv = CudaArray.empty(self.rm.shape), CudaArray.empty(self.rv.shape)
        save_mean, save_invstd = CudaArray.empty((D,)), CudaArray.empty((D,))
        if drop is None:
            lib.mc_batchnorm_train_fwd(x.ptr, self.gamma.ptr, self.beta.ptr, self.rm.ptr, self.rv.ptr, out.ptr,
                                       new_rm.ptr, new_rv.ptr, save_mean.ptr, save_invstd.ptr, N, D, self.eps,
                                       self.momentum)
        else:
            lib.mc_bn_dropout_fwd(x.ptr, self.gamma.ptr, self.beta.ptr, self.rm.ptr, self.rv.ptr, out.ptr,
                                  new_rm.ptr, new_rv.ptr, save_mean.ptr, save_invstd.ptr, N, D, self.eps,
                                  self.momentum, float(drop[0]), int(drop[1]), int(drop[2]))
        if not self.stats_done:
            self.stats_done = True
            self.save_mean, self.save_invstd = save_mean, save_invstd
            for lazy, value in ((self.new_rm, new_rm), (self.new_rv, new_rv)):
                if lazy.pending():
                    lazy.bind(value)

    def run_plain(self, _which=None):
        """Evaluate everything that is still pending with the un-fused kernel."""
        out = CudaArray.empty(self.x.shape)
        self._launch(out, None)
        if self.out.pending():
            self.out.bind(out)

    def run_with_dropout(self, prob, seed, offset):
        """The fused launch: returns dropout(batchnorm(x)); `out` itself stays unevaluated."""
        y = CudaArray.empty(self.x.shape)
        self.drop = (float(prob), int(seed), int(offset))
        self._launch(y, self.drop)
        return y


def _bn_fusable(N, D):
    return N <= 1024 and D % 8 == 0 and N * D <= (1 << 20)      # mc_bn_fused_supported


def fused_batchnorm(x, gamma, beta, running_mean, running_var, eps, momentum):
    """Train-mode batchnorm of nn/layers.py:93-110: returns (out, new_running_mean, new_running_var).
    Evaluated lazily so that a dropout called on `out` next is applied inside the same kernel."""
    lib = ops.L()
    N, D = x.shape
    if _FUSE_FORWARD and _bn_fusable(N, D) and type(x) is not LazyArray:  # pylint: disable=unidiomatic-typecheck
        pend = _BnPending(x, gamma, beta, running_mean, running_var, eps, momentum)
        pend.out = LazyArray((N, D), pend.run_plain, saved=pend)
        pend.new_rm = LazyArray(running_mean.shape, pend.run_plain)
        pend.new_rv = LazyArray(running_var.shape, pend.run_plain)
        return pend.out, pend.new_rm, pend.new_rv
    out = CudaArray.empty((N, D))
    new_rm, new_rv = CudaArray.empty(running_mean.shape), CudaArray.empty(running_var.shape)
    save_mean, save_invstd = CudaArray.empty((D,)), CudaArray.empty((D,))
    lib.mc_batchnorm_train_fwd(x.ptr, gamma.ptr, beta.ptr, running_mean.ptr, running_var.ptr, out.ptr,
                               new_rm.ptr, new_rv.ptr, save_mean.ptr, save_invstd.ptr, N, D, float(eps),
                               float(momentum))
    out._saved = (save_mean, save_invstd)  # pylint: disable=protected-access
    return out, new_rm, new_rv


def _batchnorm_vjp(ans, x, gamma, beta, running_mean, running_var, eps, momentum):  # pylint: disable=unused-argument
    out = ans[0]
    saved = out._saved  # pylint: disable=protected-access

    def vjp(gs):
        g = gs[0] if isinstance(gs, tuple) else gs
        if g is None:
            return [0.0, 0.0, 0.0]
        lib = ops.L()
        N, D = x.shape
        if isinstance(saved, _BnPending):
            if not saved.stats_done:
                saved.run_plain()
            save_mean, save_invstd = saved.save_mean, saved.save_invstd
        else:
            save_mean, save_invstd = saved
        dx, dgamma, dbeta = CudaArray.empty((N, D)), CudaArray.empty(gamma.shape), CudaArray.empty(beta.shape)
        if not _bn_fusable(N, D):
            g = concrete(g, out.shape)
            lib.mc_batchnorm_train_bwd(x.ptr, g.ptr, gamma.ptr, save_mean.ptr, save_invstd.ptr, dx.ptr, dgamma.ptr,
                                       dbeta.ptr, N, D)
            return [dx, dgamma, dbeta]
        # one kernel for: the VJP of the dropout behind this batchnorm (when the incoming gradient is exactly
        # "dy times that dropout's keep-mask": the mask is regenerated while dy is loaded), the batchnorm VJP,
        # and -- when x is the signed-zero output of a fused affine+relu -- the relu VJP plus the bias
        # gradient of that affine layer (column sums of the result)
        drop = None
        if isinstance(g, ChainGrad) and len(g.steps) == 1 and g.steps[0][0] == DROPOUT and g.g is not None \
                and g.g_mode == FULL and g.g_scalar == 1.0 and g.g.size == N * D:
            drop, g = g.steps[0], g.g
        else:
            g = concrete(g, out.shape)
        relu_src = x._tag is _RELU_SIGNED  # pylint: disable=protected-access
        dbias = CudaArray.empty((D,)) if relu_src else None
        lib.mc_bn_dropout_relu_bwd(x.ptr, g.ptr, gamma.ptr, save_mean.ptr, save_invstd.ptr, dx.ptr, dgamma.ptr, dbeta.ptr,
                                   dbias.ptr if relu_src else None, N, D, 1 if drop is not None else 0,
                                   drop[5] if drop is not None else 0.0, drop[1] if drop is not None else 0,
                                   drop[3] if drop is not None else 0, 1 if relu_src else 0)
        if relu_src:
            dx._tag = ('relu_done', x, dbias)  # pylint: disable=protected-access
        return [dx, dgamma, dbeta]
    return vjp


def fused_batchnorm_test(x, gamma, beta, running_mean, running_var, eps):
    """Test-mode batchnorm of nn/layers.py:111-113."""
    N, D = x.shape
    out = CudaArray.empty((N, D))
    ops.L().mc_batchnorm_test_fwd(x.ptr, gamma.ptr, beta.ptr, running_mean.ptr, running_var.ptr, out.ptr, N, D,
                                  float(eps))
    return out


def _batchnorm_test_vjp(ans, x, gamma, beta, running_mean, running_var, eps):  # pylint: disable=unused-argument
    """Test-mode statistics are constants: dx = g * gamma / sqrt(rv + eps), dgamma = sum(g * xhat, 0),
    dbeta = sum(g, 0).  Not on the training path (only gradient checks reach it): generic ops."""
    def vjp(g):
        lib = ops.L()
        g = concrete(g, ans.shape)
        inv = ops.unary(lib.MC_RECIP, ops.unary(lib.MC_SQRT, ops.binary(lib.MC_ADD, running_var, float(eps))))
        scale = ops.binary(lib.MC_MUL, ops.as_f32(gamma).reshape(inv.shape), inv)
        dx = ops.binary(lib.MC_MUL, g, scale)
        xhat = ops.binary(lib.MC_MUL, ops.binary(lib.MC_SUB, x, running_mean), inv)
        dgamma = ops.reduce(lib.MC_RED_SUM, ops.binary(lib.MC_MUL, g, xhat), 0).reshape(gamma.shape)
        dbeta = ops.reduce(lib.MC_RED_SUM, g, 0).reshape(beta.shape)
        return [dx, dgamma, dbeta]
    return vjp


def fused_dropout(x, prob, seed, offset):
    """x * (u > prob), not rescaled (nn/layers.py:141-144); u from Philox(seed, offset).  On a pending
    batchnorm: that batchnorm's kernel runs now and applies the mask at its store."""
    if type(x) is LazyArray and x.pending() and isinstance(x._saved, _BnPending) and not x._saved.stats_done:  # pylint: disable=protected-access,unidiomatic-typecheck
        out = x._saved.run_with_dropout(prob, seed, offset)  # pylint: disable=protected-access
        out._saved = ('after_bn',)  # pylint: disable=protected-access
        return out
    out = CudaArray.empty(x.shape)
    if x.size:
        ops.L().mc_dropout(x.ptr, out.ptr, x.size, float(prob), int(seed), int(offset))
    return out


def _dropout_vjp(ans, x, prob, seed, offset):  # pylint: disable=unused-argument
    if ans._saved == ('after_bn',):  # pylint: disable=protected-access
        # fused behind a batchnorm: leave the mask as a marker factor; that batchnorm's VJP applies it
        return lambda g: as_chain(g, ans.shape).then(*drop_step(prob, seed, offset))

    def vjp(g):
        g = concrete(g, ans.shape)
        return fused_dropout(g, prob, seed, offset)      # same (seed, offset) => same mask
    return vjp


def fused_softmax_loss(x, label):
    """softmax_loss of nn/layers.py:201-221 in one kernel: mean cross-entropy of row softmax;
    `label` is (N,) class indices or (N, C) one-hot / soft targets."""
    lib = ops.L()
    rows, cols = x.shape
    loss = CudaArray.empty(())
    if _FUSE_FORWARD and type(x) is LazyArray and x.pending() and type(x._saved) is tuple and x._saved[0] == 'affine' \
            and label.ndim == 1 and cols <= 16 and x._saved[1].shape[1] * cols * 4 <= 200 * 1024:  # pylint: disable=protected-access,unidiomatic-typecheck
        # the classifier head in one launch: the pending affine's GEMM, the loss, its gradient AND the gradient with
        # respect to the affine's input (mc_head_softmax_loss) -- three launches on the chain otherwise
        _, hh, w, b1 = x._saved  # pylint: disable=protected-access
        K = hh.shape[1]
        logits, dl, dh = CudaArray.empty((rows, cols)), CudaArray.empty((rows, cols)), CudaArray.empty((rows, K))
        lib.mc_head_softmax_loss(hh.ptr, ops.as_f32(w).ptr, b1.ptr, ops.as_i32(label).ptr, logits.ptr, dl.ptr, dh.ptr, loss.ptr,
                                 rows, K, cols)
        x.bind(logits)                                  # the affine's value, should anything else ask for it
        dl._tag = ('head_done', x, dh)  # pylint: disable=protected-access
        loss._saved = dl  # pylint: disable=protected-access
        return loss
    # the VJP ignores its upstream gradient (numpy_core.py:170-175), so dx = (p - y) / N is known now: the
    # forward kernel writes it instead of the probabilities and the backward pass launches nothing
    dx = CudaArray.empty((rows, cols))
    if label.ndim == 1:
        lib.mc_softmax_ce_fused(x.ptr, ops.as_i32(label).ptr, None, None, dx.ptr, loss.ptr, rows, cols)
    else:
        lib.mc_softmax_ce_fused(x.ptr, None, ops.as_f32(label).ptr, None, dx.ptr, loss.ptr, rows, cols)
    loss._saved = dx  # pylint: disable=protected-access
    return loss


def _softmax_loss_vjp(ans, x, label):  # pylint: disable=unused-argument
    dx = ans._saved  # pylint: disable=protected-access
    # (p - y) / N whatever the upstream gradient: the reference's softmax_output VJP discards it
    # (numpy_core.py:170-175); computed by the forward kernel
    return lambda _g: dx


def fused_temporal_softmax_loss(x, y, mask):
    """temporal_softmax_loss of nn/layers.py:444-480 -- x (N, T, V) scores, y (N, T) targets, mask (N, T) -- in
    one kernel (+ the row-loss sum): the loss and, kept for the VJP, its derivative with respect to x."""
    N, T, V = x.shape
    loss, dx = CudaArray.empty(()), CudaArray.empty((N, T, V))
    m = ops.as_f32(mask) if _is_arr(mask) else None
    ops.L().mc_temporal_softmax_loss(ops.as_f32(x).ptr, ops.as_i32(y).ptr, m.ptr if m is not None else None, dx.ptr,
                                     loss.ptr, N * T, V, N)
    loss._saved = dx  # pylint: disable=protected-access
    return loss


def _temporal_softmax_loss_vjp(ans, x, y, mask):  # pylint: disable=unused-argument
    dx = ans._saved  # pylint: disable=protected-access

    def vjp(g):
        if isinstance(g, ChainGrad):
            g = g.g_scalar if (g.g is None and not g.steps) else g.materialize()
        if not _is_arr(g):
            return ChainGrad(dx.shape, dx, float(g))          # upstream 1.0 (the usual seed): dx itself, no launch
        return ChainGrad(dx.shape, g, g_mode=ONE).then(E['MC_VJP_MUL_A'], dx, FULL)
    return vjp


def fused_lstm_step(x, prev_h, prev_c, Wx, Wh, b):
    """lstm_step of nn/layers.py:353-380: two GEMMs accumulated into one (N, 4H) buffer
    (bias in the first epilogue) and one gate kernel."""
    lib = ops.L()
    N, H = prev_c.shape
    a = ops.gemm(ops.as_f32(x), Wx, bias=b)
    ops.gemm(prev_h, Wh, out=a, accumulate=True)
    next_h, next_c, gates = CudaArray.empty((N, H)), CudaArray.empty((N, H)), CudaArray.empty((N, 4 * H))
    lib.mc_lstm_cell_fwd(a.ptr, prev_c.ptr, next_h.ptr, next_c.ptr, gates.ptr, N, H)
    next_h._saved = gates  # pylint: disable=protected-access
    return next_h, next_c


def _lstm_step_vjp(ans, x, prev_h, prev_c, Wx, Wh, b):
    next_h, next_c = ans
    gates = next_h._saved  # pylint: disable=protected-access

    def vjp(gs):
        dh, dc = gs
        N, H = prev_c.shape
        dh = concrete(dh, (N, H)) if dh is not None else ops.zeros((N, H))
        dc = concrete(dc, (N, H)) if dc is not None else None
        da, dprev_c = CudaArray.empty((N, 4 * H)), CudaArray.empty((N, H))
        ops.L().mc_lstm_cell_bwd(gates.ptr, prev_c.ptr, next_c.ptr, dh.ptr, dc.ptr if dc is not None else None,
                                 da.ptr, dprev_c.ptr, N, H)
        xf = ops.as_f32(x)
        return [ops.gemm(da, Wx, trans_b=True), ops.gemm(da, Wh, trans_b=True), dprev_c,
                ops.gemm(xf, da, trans_a=True), ops.gemm(prev_h, da, trans_a=True),
                ops.reduce(E['MC_RED_SUM'], da, 0).reshape(b.shape)]
    return vjp


def _view(base, byte_offset, shape):
    """Zero-copy view of `base`'s buffer starting `byte_offset` bytes in."""
    return CudaArray(base.buf, base.ptr + byte_offset, tuple(shape), base.dtype)


def fused_lstm_temporal(x, h0, Wx, Wh, b):
    """lstm_temporal of nn/layers.py:383-420 (x (N,T,D), h0 (N,H), zero initial cell) as ONE primitive.

    The reference slices `x[:, t, :]`, runs `lstm_step` and `append`s the new hidden state to a growing
    (N, t, H) array: O(T^2) copies forward, a scatter-add into a fresh zero tensor per slice
    backward.  Here the sequence lives TIME-MAJOR on the device, so every
    per-step operand is a contiguous slice of one big buffer and nothing is sliced or appended:
      * the input projection of all T steps is one GEMM, x_tm (T*N, D) . Wx + b;
      * step t is one accumulate-GEMM  a_t += h_{t-1} . Wh  and one gate kernel that writes h_t / c_t
        straight into slot t+1 of the (T+1, N, H) state buffers and the gates into (T, N, 4H);
      * the result is one transpose back to (N, T, H).
    Same arithmetic per element as T calls of `fused_lstm_step` (gate order i, f, o, g)."""
    lib = ops.L()
    x, h0 = ops.as_f32(x), ops.as_f32(h0)
    N, T, D = x.shape
    H = h0.shape[1]
    prec = ops._gemm_precision['mode']  # pylint: disable=protected-access
    x_tm = ops.transpose(x, (1, 0, 2)).reshape((T * N, D))
    A = ops.gemm(x_tm, Wx, bias=b)                                   # (T*N, 4H), rows of step t contiguous
    hbuf, cbuf = CudaArray.empty((T + 1, N, H)), CudaArray.empty((T + 1, N, H))
    gates = CudaArray.empty((T, N, 4 * H))
    lib.mc_memcpy_d2d(hbuf.ptr, h0.ptr, 4 * N * H)
    lib.mc_memset(cbuf.ptr, 0, 4 * N * H)
    sH, sA = 4 * N * H, 4 * N * 4 * H
    # the loop over t runs in the library (csrc/lstm.cu): per step one K-sliced GEMM h_{t-1}.Wh whose
    # partials the gate kernel adds while reading a_t -- two launches, no split-K combine pass
    lib.mc_lstm_seq_fwd(A.ptr, Wh.ptr, hbuf.ptr, cbuf.ptr, gates.ptr, T, N, H, prec)
    hs = ops.transpose(_view(hbuf, sH, (T, N, H)), (1, 0, 2))
    hs._saved = (x_tm, hbuf, cbuf, gates)  # pylint: disable=protected-access
    return hs


def _lstm_temporal_vjp(ans, x, h0, Wx, Wh, b):
    x_tm, hbuf, cbuf, gates = ans._saved  # pylint: disable=protected-access

    def vjp(g):
        """BPTT over the time-major buffers: per step one gate-backward kernel and one
        accumulate-GEMM dH[t] += da_t . Wh^T (dH[t] already holds the upstream gradient of h_{t-1},
        so no separate add); the weight gradients are three reductions over ALL steps at once."""
        lib = ops.L()
        N, T, _ = x.shape
        H = h0.shape[1]
        D = x.shape[2]
        prec = ops._gemm_precision['mode']  # pylint: disable=protected-access
        g = concrete(g, ans.shape)
        sH, sA = 4 * N * H, 4 * N * 4 * H
        dH = CudaArray.empty((T + 1, N, H))
        lib.mc_memset(dH.ptr, 0, sH)
        ops.transpose(g, (1, 0, 2), out=_view(dH, sH, (T, N, H)))      # time-major, straight into slots 1..T
        dA = CudaArray.empty((T * N, 4 * H))
        lib.mc_lstm_seq_bwd(gates.ptr, cbuf.ptr, Wh.ptr, dH.ptr, dA.ptr, T, N, H, prec)
        h_prev = _view(hbuf, 0, (T * N, H))
        # (the three parameter gradients ran on the side stream beside dx until r2x: all four contractions stream the
        #  210 MB of dA and fill every SM, and A/B on one box gave 2.439 ms/step beside each other, 2.399 in sequence)
        dWx = ops.gemm(x_tm, dA, trans_a=True)
        dWh = ops.gemm(h_prev, dA, trans_a=True)
        db = ops.reduce(E['MC_RED_SUM'], dA, 0).reshape(b.shape)
        dh0 = CudaArray.empty((N, H))
        lib.mc_memcpy_d2d(dh0.ptr, dH.ptr, sH)
        dx = ops.transpose(ops.gemm(dA, Wx, trans_b=True).reshape((T, N, D)), (1, 0, 2))
        return [dx, dh0, dWx, dWh, db]
    return vjp


# ---- optimizer updates (nn/optim.py:33-141), one kernel per parameter tensor ------------
def _grad_like(w, dw):
    """The gradient as a float32 device array of w's shape.  A parameter the loss does not reach
    gets the Python scalar 0.0 from the tape (tape.py:249-257); the reference's NumPy arithmetic
    broadcasts it, so do the kernels' callers."""
    if ops.is_scalar(dw):
        return ops.full(w.shape, float(dw))
    return ops.as_f32(dw)


def fused_sgd(w, dw, lr):
    out = CudaArray.empty(w.shape)
    ops.L().mc_sgd(w.ptr, _grad_like(w, dw).ptr, out.ptr, float(lr), w.size)
    return out


def fused_sgd_momentum(w, dw, v, lr, momentum):
    w_out, v_out = CudaArray.empty(w.shape), CudaArray.empty(w.shape)
    ops.L().mc_sgd_momentum(w.ptr, _grad_like(w, dw).ptr, v.ptr, w_out.ptr, v_out.ptr, float(lr), float(momentum), w.size)
    return w_out, v_out


def fused_rmsprop(w, dw, cache, lr, decay, eps):
    w_out, c_out = CudaArray.empty(w.shape), CudaArray.empty(w.shape)
    ops.L().mc_rmsprop(w.ptr, _grad_like(w, dw).ptr, cache.ptr, w_out.ptr, c_out.ptr, float(lr), float(decay), float(eps),
                       w.size)
    return w_out, c_out


def fused_adam(w, dw, m, v, lr, beta1, beta2, eps, t):
    w_out, m_out, v_out = CudaArray.empty(w.shape), CudaArray.empty(w.shape), CudaArray.empty(w.shape)
    ops.L().mc_adam(w.ptr, _grad_like(w, dw).ptr, m.ptr, v.ptr, w_out.ptr, m_out.ptr, v_out.ptr, float(lr), float(beta1),
                    float(beta2), float(eps), int(t), w.size)
    return w_out, m_out, v_out


# ======================================================================================
#  registration
# ======================================================================================
def register_primitives(reg, prim_wrapper):
    for name, (opcode, _, _) in _UNARY.items():
        reg.register(name, prim_wrapper(_make_unary(name, opcode)))
    for name, opcode in list(_BINARY.items()) + list(_COMPARE.items()):
        reg.register(name, prim_wrapper(_make_binary(name, opcode)))
    reg.register('square', prim_wrapper(square))
    for name, opcode in (('sum', 'MC_RED_SUM'), ('max', 'MC_RED_MAX'), ('amax', 'MC_RED_MAX'),
                         ('min', 'MC_RED_MIN'), ('amin', 'MC_RED_MIN')):
        reg.register(name, prim_wrapper(_reduce_fn(name, opcode)))
    for fn in (argmax, mean, count_nonzero, reshape, transpose, expand_dims, squeeze, ravel, broadcast_to,
               append, concatenate, stack, _minpy_getitem, zeros, ones, empty, full, zeros_like, ones_like,
               array, arange, eye, dot, softmax_output, onehot_encode,
               fused_affine, fused_relu, fused_batchnorm, fused_batchnorm_test, fused_dropout,
               fused_softmax_loss, fused_temporal_softmax_loss, fused_lstm_step, fused_lstm_temporal, fused_sgd, fused_sgd_momentum, fused_rmsprop, fused_adam):
        reg.register(fn.__name__, prim_wrapper(fn))


def def_grads(prims):
    for name, (_, vjp_op, operand) in _UNARY.items():
        if vjp_op is None:
            if name in ('sign', 'round', 'ceil', 'floor'):
                prims(name).def_grad_zero()
            continue

        def maker(ans, x, _op=vjp_op, _operand=operand):
            src = {'x': x, 'ans': ans, None: None}[_operand]
            return lambda g: as_chain(g, ans.shape).then(E[_op], src, FULL)
        prims(name).def_grad(maker)
    prims('square').def_grad(
        lambda ans, x: lambda g: as_chain(g, ans.shape).then(E['MC_VJP_MUL_A'], x, FULL).then(
            E['MC_VJP_MUL_S'], scalar=2.0))
    for name in _BINARY:
        dx, dy = _binary_vjps(name)
        prims(name).def_grad(dx)
        prims(name).def_grad(dy, argnum=1)

    prims('sum').def_grad(_sum_vjp)
    for name in ('max', 'amax', 'min', 'amin'):
        prims(name).def_grad(_extremum_vjp)
    prims('mean').def_grad(lambda ans, x, axis=None, keepdims=False: (
        lambda g: as_chain(_sum_vjp(ans, x, axis, None, keepdims)(g), x.shape).then(
            E['MC_VJP_MUL_S'], scalar=float(max(1, ans.size)) / max(1, x.size))))

    for name in ('reshape', 'expand_dims', 'squeeze', 'ravel'):
        prims(name).def_grad(_reshape_back)
    prims('transpose').def_grad(_transpose_vjp)
    prims('broadcast_to').def_grad(
        lambda ans, x, shape: lambda g: ops.unbroadcast(concrete(g, ans.shape), x.shape))
    prims('append').def_grad(_append_vjp(0))
    prims('append').def_grad(_append_vjp(1), argnum=1)
    prims('_minpy_getitem').def_grad(_getitem_vjp)

    prims('dot').def_grad(_dot_vjp_a)
    prims('dot').def_grad(_dot_vjp_b, argnum=1)
    prims('softmax_output').def_grad(_softmax_output_vjp)

    dx, dw, db = _affine_vjps()
    prims('fused_affine').def_grad(dx).def_grad(dw, argnum=1).def_grad(db, argnum=2)
    prims('fused_relu').def_grad(_relu_vjp)
    prims('fused_batchnorm').def_multiple_grad(_batchnorm_vjp, (0, 1, 2))
    prims('fused_batchnorm').def_grad_zero(3).def_grad_zero(4)
    prims('fused_batchnorm_test').def_multiple_grad(_batchnorm_test_vjp, (0, 1, 2))
    prims('fused_batchnorm_test').def_grad_zero(3).def_grad_zero(4)
    prims('fused_dropout').def_grad(_dropout_vjp)
    prims('fused_softmax_loss').def_grad(_softmax_loss_vjp)
    prims('fused_temporal_softmax_loss').def_grad(_temporal_softmax_loss_vjp).def_grad_zero(1).def_grad_zero(2)
    prims('fused_lstm_step').def_multiple_grad(_lstm_step_vjp, (0, 1, 2, 3, 4, 5))
    prims('fused_lstm_temporal').def_multiple_grad(_lstm_temporal_vjp, (0, 1, 2, 3, 4))
