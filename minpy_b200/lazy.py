"""ChainGrad: a gradient that is still a recipe -- `g * f1 * f2 * ...` over one shape.

This is the host half of the tape rewrite the north star asks for.  In the reference every
VJP closure launches its own array ops (minpy/tape.py:223-234 calls one closure per record;
e.g. numpy_core.py:240-243 `g * y`).  Here the VJPs of elementwise CUDA primitives do not
launch anything: they append one `mc_vjp_step` to the incoming gradient's recipe.  The recipe
turns into ONE `mc_vjp_chain` kernel launch when something needs real numbers: a GEMM /
reduction VJP, an accumulation into a gradient that already exists (fused into the same
kernel through `acc`), a shape change, or the end of the backward pass.
"""
from __future__ import annotations

import numpy as np

from . import ops
from .cuda_array import CudaArray, _Buffer, _state as _cuda_state

FULL, ROW, COL, ONE = 0, 1, 2, 3
DROPOUT = 'DROPOUT'     # marker step: the keep-mask of a dropout, regenerated from (seed, offset); see drop_step

_PTR = CudaArray.__dict__['ptr']
_BUF = CudaArray.__dict__['buf']


class LazyArray(CudaArray):
    """A forward value whose producing kernel has NOT been launched yet (the forward counterpart of
    ChainGrad).  `layers.affine` and `layers.batchnorm` return one: if the very next layer call is one the
    producer can absorb -- relu into the GEMM epilogue, dropout into the batchnorm store -- that call
    launches the fused kernel and this object is never evaluated; anything else that touches `.ptr` /
    `.buf` (any kernel argument, asnumpy, a reshape) evaluates it on the spot through `thunk(self)`,
    which must end in `bind`.  Shape, dtype and device are known without evaluation."""
    __slots__ = ('_thunk', '_pool', '_dev')

    def __init__(self, shape, thunk, saved=None):  # pylint: disable=super-init-not-called
        self.shape = tuple(int(s) for s in shape)
        self.size = ops._prod(self.shape)  # pylint: disable=protected-access
        self.dtype = ops._F32  # pylint: disable=protected-access
        self._saved = saved
        self._tag = None
        self._readback = None
        self._thunk = thunk
        self._pool = _Buffer._active_pool  # pylint: disable=protected-access
        self._dev = _cuda_state['device']

    def pending(self):
        return self._thunk is not None

    def bind(self, arr):
        """Adopt the storage of the computed array."""
        _BUF.__set__(self, arr.buf)
        _PTR.__set__(self, arr.ptr)
        self._thunk = None

    def _force(self):
        thunk = self._thunk
        if thunk is not None:
            if self._pool is not _Buffer._active_pool:  # pylint: disable=protected-access
                raise RuntimeError('a lazily evaluated value of a captured step was used outside that capture')
            thunk(self)
            if self._thunk is not None:
                raise RuntimeError('lazy value: the thunk did not bind a result')

    @property
    def ptr(self):
        if self._thunk is not None:
            self._force()
        return _PTR.__get__(self, CudaArray)

    @property
    def buf(self):
        if self._thunk is not None:
            self._force()
        return _BUF.__get__(self, CudaArray)

    @property
    def device(self):
        return self._dev


def drop_step(prob, seed, offset):
    """Marker factor of a ChainGrad: multiply by the keep-mask of dropout(prob, seed, offset).  Never reaches
    mc_vjp_chain: batchnorm's VJP consumes it inside its fused kernel; `materialize` otherwise runs
    mc_dropout between the chain segments."""
    return (DROPOUT, int(seed), FULL, int(offset), FULL, float(prob), 1)


class ChainGrad(object):
    """`out[i] = g[gmap(i)] * g_scalar * prod_k step_k(i)` over a dense array of `shape`."""
    __slots__ = ('shape', 'g', 'g_scalar', 'g_mode', 'g_inner', 'steps')

    def __init__(self, shape, g=None, g_scalar=1.0, g_mode=FULL, g_inner=1, steps=()):
        self.shape = tuple(shape)
        self.g = g
        self.g_scalar = float(g_scalar)
        self.g_mode = g_mode
        self.g_inner = g_inner
        self.steps = tuple(steps)

    @property
    def size(self):
        return ops._prod(self.shape)

    @property
    def ndim(self):
        return len(self.shape)

    def then(self, op, a=None, a_mode=FULL, b=None, b_mode=FULL, scalar=0.0, inner=1):
        """A new recipe with one more factor (the receiver is unchanged: it may fan out)."""
        step = (op, a, a_mode, b, b_mode, float(scalar), int(inner))
        return ChainGrad(self.shape, self.g, self.g_scalar, self.g_mode, self.g_inner, self.steps + (step,))

    def reshaped(self, shape):
        """Same elements under another shape; only valid while every operand is FULL/ONE."""
        shape = tuple(shape)
        if self.g_mode in (FULL, ONE) and all(s[2] in (FULL, ONE) and s[4] in (FULL, ONE) for s in self.steps):
            g = self.g
            if g is not None and self.g_mode == FULL:
                g = g.reshape(shape)
            return ChainGrad(shape, g, self.g_scalar, self.g_mode, self.g_inner, self.steps)
        return ChainGrad(shape, self.materialize().reshape(shape))

    def _materialize_mixed(self, acc):
        """A recipe that contains dropout marker steps: chain segment, mc_dropout, chain segment, ..."""
        cur = ChainGrad(self.shape, self.g, self.g_scalar, self.g_mode, self.g_inner)
        pending = []
        for step in self.steps:
            if step[0] != DROPOUT:
                pending.append(step)
                continue
            base = ChainGrad(cur.shape, cur.g, cur.g_scalar, cur.g_mode, cur.g_inner, tuple(pending)).materialize() \
                if (pending or cur.g is None or cur.g_mode != FULL or cur.g_scalar != 1.0) else cur.g
            if base.shape != self.shape:
                base = base.reshape(self.shape)
            out = CudaArray.empty(self.shape)
            if out.size:
                ops.L().mc_dropout(base.ptr, out.ptr, out.size, step[5], step[1], step[3])
            cur, pending = ChainGrad(self.shape, out), []
        return ChainGrad(self.shape, cur.g, cur.g_scalar, cur.g_mode, cur.g_inner, tuple(pending)).materialize(acc)

    def materialize(self, acc=None):
        """Run the recipe: one kernel per 12 steps; `acc` (same shape) is added in the same
        kernel, which is how gradient accumulation on the tape is fused."""
        for step in self.steps:
            if step[0] == DROPOUT:
                return self._materialize_mixed(acc)
        if not self.steps and acc is None and self.g is not None and self.g_mode == FULL and self.g_scalar == 1.0:
            return self.g if self.g.shape == self.shape else self.g.reshape(self.shape)
        return ops.run_chain(self.g, self.g_scalar, list(self.steps), self.shape, acc=acc,
                             g_mode=self.g_mode, g_inner=self.g_inner)


def as_chain(g, shape):
    """View any incoming gradient (recipe, device array, host array, Python scalar) as a
    recipe over `shape`; size-1 gradients broadcast."""
    shape = tuple(shape)
    if isinstance(g, ChainGrad):
        if g.shape == shape:
            return g
        if g.size == ops._prod(shape):
            return g.reshaped(shape)
        g = g.materialize()
    if isinstance(g, np.ndarray):
        if g.ndim == 0:
            return ChainGrad(shape, None, float(g))
        g = CudaArray.from_numpy(g)
    if isinstance(g, CudaArray):
        if g.dtype != ops._F32:
            g = g.astype(ops._F32)
        if g.shape == shape:
            return ChainGrad(shape, g)
        if g.size == ops._prod(shape):
            return ChainGrad(shape, g.reshape(shape))
        if g.size == 1:
            return ChainGrad(shape, g, g_mode=ONE)
        raise ValueError('gradient of shape %s does not match result shape %s' % (g.shape, shape))
    return ChainGrad(shape, None, float(g))


def concrete(g, shape=None):
    """Device array for a gradient in any form (launches the chain kernel if needed)."""
    if isinstance(g, ChainGrad):
        return g.materialize()
    if isinstance(g, CudaArray):
        return g if g.dtype == ops._F32 else g.astype(ops._F32)
    if isinstance(g, np.ndarray) and g.ndim > 0:
        return CudaArray.from_numpy(g).astype(ops._F32)
    if shape is None:
        raise TypeError('cannot make a device array from scalar gradient without a shape')
    return ops.full(shape, float(g))
