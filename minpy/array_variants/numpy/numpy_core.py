"""Primitives and VJPs of the NumPy array variant.

Restates, in table form, what the reference defines in
minpy/array_variants/numpy/numpy_core.py:211-304: every routine/ufunc of the `numpy`
namespace becomes a primitive, four extras are added (`_minpy_getitem`, `sigmoid`,
`onehot_encode`, `softmax_output`), and about forty-five of them get vector-Jacobian
products.  The arithmetic of each VJP is dictated by the reference (cited per entry) because
this variant is the oracle the CUDA variant is judged against.
"""
from __future__ import absolute_import
from __future__ import division

import inspect

import numpy as np


# ------------------------------------------------------------------ extra primitives ----
def _minpy_getitem(arr, index):
    return arr[index]


def _sigmoid(x):
    """Logistic function evaluated without overflow on either tail (numpy_core.py:32-43)."""
    e = np.exp(-np.abs(x))
    return np.where(x >= 0, 1.0, e) / (1.0 + e)


def _onehot_encode(indices, out):
    """out[i, indices[i]] = 1, in place; returns `out` (numpy_core.py:47-63)."""
    out[np.arange(indices.shape[0]), indices] = 1
    return out


def _softmax_output(x, _label):
    """Row softmax; the label only matters to the VJP (numpy_core.py:164-168)."""
    shifted = np.exp(x - np.max(x, axis=1, keepdims=True))
    return shifted / np.sum(shifted, axis=1, keepdims=True)


# ------------------------------------------------------------------ VJP building blocks --
def _sum_to_shape(grad, shape):
    """Undo NumPy broadcasting: sum leading axes away, then every size-1 axis with keepdims
    (numpy_core.py:119-146)."""
    while np.ndim(grad) > len(shape):
        grad = np.sum(grad, axis=0)
    for ax, extent in enumerate(shape):
        if extent == 1:
            grad = np.sum(grad, axis=ax, keepdims=True)
    return grad


def _bcast_vjp(local):
    """Lift `local(g, ans, x, y)` (gradient in the broadcast shape) to a VJP maker for the
    operand at `pos`, folding back to that operand's own shape."""
    def maker(pos):
        def vjp_maker(ans, x, y):
            operand = (x, y)[pos]
            if isinstance(operand, np.ndarray):
                shape = operand.shape
                return lambda g: _sum_to_shape(local(g, ans, x, y), shape)
            if isinstance(ans, np.ndarray):          # scalar operand, array result
                return lambda g: np.sum(local(g, ans, x, y))
            return lambda g: local(g, ans, x, y)     # scalar op scalar
        return vjp_maker
    return maker


def _reduction_axes(axis, ndim):
    if axis is None:
        return tuple(range(ndim))
    if isinstance(axis, (int, np.integer)):
        return (int(axis) % ndim,)
    return tuple(int(a) % ndim for a in axis)


def _expander(x, axis):
    """Function re-inflating a reduced result/gradient to x's shape (as float64 zeros + g,
    which is what makes the reference's sum-gradients float64: numpy_core.py:150-162)."""
    if not isinstance(x, np.ndarray):
        return lambda g: g
    xshape = x.shape
    axes = _reduction_axes(axis, len(xshape))
    kept = tuple(1 if i in axes else n for i, n in enumerate(xshape))
    if axis is None:
        return lambda g: np.full(xshape, g)
    return lambda g: np.zeros(xshape) + np.reshape(g, kept)


def _sum_vjp(ans, x, axis=None, keepdims=False):  # pylint: disable=unused-argument
    return _expander(x, axis)


def _extremum_vjp(ans, x, axis=None, out=None, keepdims=False):  # pylint: disable=unused-argument
    """max/min: every position equal to the extremum receives the whole gradient -- ties are
    not split (numpy_core.py:66-70; pinned by tests/unittest/test_autograd.py:107-130)."""
    expand = _expander(x, axis)
    if isinstance(x, np.ndarray):
        if axis is None:
            hit = (x == ans)
        else:
            axes = _reduction_axes(axis, x.ndim)
            kept = tuple(1 if i in axes else n for i, n in enumerate(x.shape))
            hit = (x == np.reshape(ans, kept))
    else:
        hit = 1.0
    return lambda g: expand(g) * hit


def _getitem_vjp(ans, x, index):  # pylint: disable=unused-argument
    def vjp(g):
        full = np.zeros_like(x)
        np.add.at(full, index, g)       # duplicates accumulate (numpy_core.py:24-28)
        return full
    return vjp


def _append_vjp(pos):
    def maker(ans, arr, values, axis=None):  # pylint: disable=unused-argument
        return lambda g: np.split(g, [arr.shape[axis]], axis)[pos]
    return maker


def _softmax_output_vjp(ans, x, y):
    """(p - y) / N and nothing else: the upstream gradient is ignored by design
    (numpy_core.py:170-175)."""
    return lambda _g: (ans - y) / x.shape[0]


# Broadcasting binary ufuncs: name -> (d/dx, d/dy) as functions of (g, ans, x, y).
_BINARY_VJPS = {
    'add':         (lambda g, ans, x, y: g,                 lambda g, ans, x, y: g),
    'subtract':    (lambda g, ans, x, y: g,                 lambda g, ans, x, y: -g),
    'multiply':    (lambda g, ans, x, y: g * y,             lambda g, ans, x, y: x * g),
    'divide':      (lambda g, ans, x, y: g / y,             lambda g, ans, x, y: -g * x / y ** 2),
    'true_divide': (lambda g, ans, x, y: g / y,             lambda g, ans, x, y: -g * x / y ** 2),
    'mod':         (lambda g, ans, x, y: g,                 lambda g, ans, x, y: -g * np.floor(x / y)),
    # NaN at x == 0, as in the reference (numpy_core.py:277-278)
    'power':       (lambda g, ans, x, y: g * y * ans / x,   lambda g, ans, x, y: g * np.log(x) * ans),
    # equal operands both receive g (numpy_core.py:281-291)
    'maximum':     (lambda g, ans, x, y: g * (x == ans),    lambda g, ans, x, y: g * (y == ans)),
    'minimum':     (lambda g, ans, x, y: g * (x == ans),    lambda g, ans, x, y: g * (y == ans)),
}

# Unary elementwise: name -> function of (g, ans, x).
_UNARY_VJPS = {
    'negative': lambda g, ans, x: -g,
    'abs':      lambda g, ans, x: np.sign(x) * g,
    'absolute': lambda g, ans, x: np.sign(x) * g,
    'sqrt':     lambda g, ans, x: g * 0.5 / np.sqrt(x),
    'exp':      lambda g, ans, x: ans * g,
    'log':      lambda g, ans, x: g / x,
    'sin':      lambda g, ans, x: g * np.cos(x),
    'cos':      lambda g, ans, x: -g * np.sin(x),
    'tanh':     lambda g, ans, x: g / np.cosh(x) ** 2,
    'sigmoid':  lambda g, ans, x: g * ans * (1 - ans),
}

_ZERO_VJPS = ('sign', 'round', 'ceil', 'floor')


def register_primitives(reg, prim_wrapper):
    """Every routine / ufunc of `numpy` plus the four MinPy extras."""
    for name, obj in vars(np).items():
        if isinstance(obj, np.ufunc) or inspect.isroutine(obj):
            reg.register(name, prim_wrapper(obj))
    reg.register('_minpy_getitem', prim_wrapper(_minpy_getitem))
    reg.register('sigmoid', prim_wrapper(_sigmoid))
    reg.register('onehot_encode', prim_wrapper(_onehot_encode))
    reg.register('softmax_output', prim_wrapper(_softmax_output))


def def_grads(prims):
    """Attach the VJPs (numpy_core.py:221-304)."""
    prims('dot').def_grad(lambda ans, a, b: lambda g: np.dot(g, b.T))
    prims('dot').def_grad(lambda ans, a, b: lambda g: np.dot(a.T, g), argnum=1)

    for name, (dx, dy) in _BINARY_VJPS.items():
        prims(name).def_grad(_bcast_vjp(dx)(0))
        prims(name).def_grad(_bcast_vjp(dy)(1), argnum=1)
    for name, local in _UNARY_VJPS.items():
        prims(name).def_grad(
            (lambda f: lambda ans, x: lambda g: f(g, ans, x))(local))
    for name in _ZERO_VJPS:
        prims(name).def_grad_zero()

    prims('sum').def_grad(_sum_vjp)
    for name in ('max', 'amax', 'min', 'amin'):
        prims(name).def_grad(_extremum_vjp)

    prims('broadcast_to').def_grad(
        lambda ans, x, shape: (lambda g: _sum_to_shape(g, np.shape(x))))
    prims('transpose').def_grad(lambda ans, x, axes=None: (
        (lambda g: np.transpose(g)) if axes is None else (lambda g: np.transpose(g, np.argsort(axes)))))
    prims('reshape').def_grad(lambda ans, x, _shape: lambda g: np.reshape(g, np.shape(x)))
    prims('expand_dims').def_grad(lambda ans, x, axis: lambda g: np.reshape(g, np.shape(x)))
    prims('append').def_grad(_append_vjp(0))
    prims('append').def_grad(_append_vjp(1), argnum=1)
    prims('_minpy_getitem').def_grad(_getitem_vjp)
    prims('softmax_output').def_grad(_softmax_output_vjp)
