"""Functional operators on CudaArray -- thin host glue over the C ABI.

Each function picks the kernel variant (same-shape / scalar / vector-broadcast / general
broadcast), allocates the result from the stream-ordered pool and enqueues ONE library
call; nothing here synchronises.  The `cuda` array variant
(minpy/array_variants/cuda/cuda_core.py) registers these as MinPy primitives.
"""
from __future__ import annotations

import ctypes
import numbers

import numpy as np

from ._lib import get_lib, VjpStep
from .cuda_array import CudaArray, _F32, _I32, _I64, _prod, ensure_device

_lib = None


def L():
    global _lib
    if _lib is None:
        _lib = get_lib()
        ensure_device()
    return _lib


def is_scalar(x):
    return isinstance(x, (numbers.Number, np.generic)) or (isinstance(x, np.ndarray) and x.ndim == 0)


def as_cuda(x):
    """Coerce an operand to CudaArray (host arrays are uploaded)."""
    if isinstance(x, CudaArray):
        return x
    if isinstance(x, np.ndarray):
        return CudaArray.from_numpy(x)
    if isinstance(x, (list, tuple)):
        return CudaArray.from_numpy(np.asarray(x))
    raise TypeError('cannot convert %r to CudaArray' % type(x))


def as_f32(x):
    x = as_cuda(x)
    return x if x.dtype == _F32 else x.astype(_F32)


def as_i32(x):
    x = as_cuda(x)
    return x if x.dtype == _I32 else x.astype(_I32)


def empty(shape, dtype=_F32):
    return CudaArray.empty(shape, dtype)


def full(shape, value, dtype=None):
    if isinstance(shape, (int, np.integer)):
        shape = (int(shape),)
    out = CudaArray.empty(tuple(shape))
    if out.size:
        L().mc_fill(out.ptr, float(value), out.size)
    return out


def zeros(shape, dtype=None):
    if isinstance(shape, (int, np.integer)):
        shape = (int(shape),)
    out = CudaArray.empty(tuple(shape))
    if out.size:
        L().mc_memset(out.ptr, 0, out.nbytes)
    return out


def ones(shape, dtype=None):
    return full(shape, 1.0)


def zeros_like(x, dtype=None):
    return zeros(x.shape)


def ones_like(x, dtype=None):
    return full(x.shape, 1.0)


def arange(start, stop=None, step=1, dtype=None):
    if stop is None:
        start, stop = 0, start
    n = max(0, int(np.ceil((stop - start) / float(step))))
    integral = all(isinstance(v, (int, np.integer)) for v in (start, stop, step))
    out = CudaArray.empty((n,))
    if n:
        L().mc_arange(out.ptr, float(start), float(step), n)
    if integral and dtype is None or (dtype is not None and np.dtype(dtype).kind in 'iu'):
        out = out.astype(_I32)
    if start == 0 and step == 1:
        out._tag = ('arange', n)
    return out


# ------------------------------------------------------------------ unary ----------
def unary(op, x):
    x = as_f32(x)
    out = CudaArray.empty(x.shape)
    if x.size:
        L().mc_unary(op, x.ptr, out.ptr, x.size)
    return out


# ------------------------------------------------------------------ binary ---------
def broadcast_shape(sa, sb):
    n = max(len(sa), len(sb))
    pa = (1,) * (n - len(sa)) + tuple(sa)
    pb = (1,) * (n - len(sb)) + tuple(sb)
    out = []
    for x, y in zip(pa, pb):
        if x == y or y == 1:
            out.append(x)
        elif x == 1:
            out.append(y)
        else:
            raise ValueError('operands could not be broadcast together with shapes %s %s' % (sa, sb))
    return tuple(out)


def _contig_strides(shape):
    st, acc = [], 1
    for s in reversed(shape):
        st.append(acc)
        acc *= s
    return tuple(reversed(st))


def _bcast_strides(shape, out_shape):
    pad = len(out_shape) - len(shape)
    st = _contig_strides(shape)
    return tuple(0 if (i < pad or shape[i - pad] == 1) else st[i - pad] for i in range(len(out_shape)))


def _collapse(shape, sa, sb):
    """Merge adjacent dims that are contiguous for both operands (keeps ndim <= 6 in practice)."""
    dims = [(n, a, b) for n, a, b in zip(shape, sa, sb) if n != 1]
    if not dims:
        return (1,), (0,), (0,)
    out = [dims[0]]
    for n, a, b in dims[1:]:
        pn, pa, pb = out[-1]
        if pa == a * n and pb == b * n:
            out[-1] = (pn * n, a, b)
        else:
            out.append((n, a, b))
    return tuple(d[0] for d in out), tuple(d[1] for d in out), tuple(d[2] for d in out)


def _vec_pattern(big, small):
    """Classify `small` broadcast against `big` (big.shape == result shape).
    Returns ('row', rows, cols) / ('col', rows, cols) / None."""
    bs, ss = big.shape, small.shape
    if len(ss) > len(bs) or not bs:
        return None
    ps = (1,) * (len(bs) - len(ss)) + tuple(ss)
    # trailing block of `small` equals trailing block of big, everything before is 1 -> row vector
    k = len(bs)
    while k > 0 and ps[k - 1] == bs[k - 1]:
        k -= 1
    if all(d == 1 for d in ps[:k]):
        cols = _prod(bs[k:])
        if cols and small.size == cols:
            return ('row', big.size // cols, cols)
    # leading block equal, everything after is 1 -> column vector
    k = 0
    while k < len(bs) and ps[k] == bs[k]:
        k += 1
    if all(d == 1 for d in ps[k:]):
        rows = _prod(bs[:k])
        if rows and small.size == rows:
            return ('col', rows, big.size // rows)
    return None


def binary(op, a, b):
    """NumPy-broadcasting binary ufunc / comparison on float32."""
    lib = L()
    a_s, b_s = is_scalar(a), is_scalar(b)
    if a_s and b_s:
        raise TypeError('binary(): at least one operand must be an array')
    if b_s:
        a = as_f32(a)
        out = CudaArray.empty(a.shape)
        if a.size:
            lib.mc_binary_scalar(op, a.ptr, float(b), 0, out.ptr, a.size)
        return out
    if a_s:
        b = as_f32(b)
        out = CudaArray.empty(b.shape)
        if b.size:
            lib.mc_binary_scalar(op, b.ptr, float(a), 1, out.ptr, b.size)
        return out
    a, b = as_f32(a), as_f32(b)
    if a.shape == b.shape:
        out = CudaArray.empty(a.shape)
        if a.size:
            lib.mc_binary(op, a.ptr, b.ptr, out.ptr, a.size)
        return out
    oshape = broadcast_shape(a.shape, b.shape)
    out = CudaArray.empty(oshape)
    if out.size == 0:
        return out
    if a.shape == oshape or a.size == out.size:
        pat = _vec_pattern(out, b)
        if pat is not None:
            lib.mc_binary_vec(op, a.ptr, b.ptr, 1 if pat[0] == 'col' else 0, 0, out.ptr, pat[1], pat[2])
            return out
    if b.shape == oshape or b.size == out.size:
        pat = _vec_pattern(out, a)
        if pat is not None:
            lib.mc_binary_vec(op, b.ptr, a.ptr, 1 if pat[0] == 'col' else 0, 1, out.ptr, pat[1], pat[2])
            return out
    shape, sa, sb = _collapse(oshape, _bcast_strides(a.shape, oshape), _bcast_strides(b.shape, oshape))
    nd = len(shape)
    if nd > 6:
        raise ValueError('broadcast over more than 6 non-mergeable dims is not supported')
    arr = ctypes.c_int64 * nd
    lib.mc_binary_bcast(op, a.ptr, b.ptr, out.ptr, nd, arr(*shape), arr(*sa), arr(*sb))
    return out


def broadcast_to(x, shape):
    x = as_f32(x)
    shape = tuple(int(s) for s in shape)
    if x.shape == shape:
        return x
    return binary(L().MC_ADD, zeros(shape), x)


# ------------------------------------------------------------------ reductions -----
def _norm_axes(axis, ndim):
    if axis is None:
        return tuple(range(ndim))
    if isinstance(axis, (int, np.integer)):
        axis = (int(axis),)
    axes = tuple(sorted(a + ndim if a < 0 else a for a in axis))
    for a in axes:
        if a < 0 or a >= ndim:
            raise ValueError('axis %d out of bounds for array of dimension %d' % (a, ndim))
    return axes


def _reduced_shape(shape, axes, keepdims):
    if keepdims:
        return tuple(1 if i in axes else s for i, s in enumerate(shape))
    return tuple(s for i, s in enumerate(shape) if i not in axes)


def _runs(axes):
    """Split sorted axes into runs of consecutive axes."""
    runs, cur = [], [axes[0]]
    for a in axes[1:]:
        if a == cur[-1] + 1:
            cur.append(a)
        else:
            runs.append(cur)
            cur = [a]
    runs.append(cur)
    return runs


def reduce(op, x, axis=None, keepdims=False):
    """sum / max / min over `axis` (None, int or tuple)."""
    lib = L()
    x = as_f32(x)
    axes = _norm_axes(axis, x.ndim)
    final_shape = _reduced_shape(x.shape, axes, keepdims)
    if not axes:
        return x
    cur, cur_shape = x, list(x.shape)
    # reduce each run of consecutive axes with one (outer, red, inner) call, last run first
    for run in reversed(_runs(list(axes))):
        outer = _prod(cur_shape[:run[0]])
        red = _prod(cur_shape[run[0]:run[-1] + 1])
        inner = _prod(cur_shape[run[-1] + 1:])
        out = CudaArray.empty((outer, inner))
        lib.mc_reduce(op, cur.ptr, out.ptr, outer, red, inner)
        cur_shape = cur_shape[:run[0]] + cur_shape[run[-1] + 1:]
        cur = out
    return cur.reshape(final_shape)


def argmax(x, axis=None):
    lib = L()
    x = as_f32(x)
    if axis is None:
        outer, red, inner, oshape = 1, x.size, 1, ()
    else:
        ax = axis + x.ndim if axis < 0 else axis
        outer, red, inner = _prod(x.shape[:ax]), x.shape[ax], _prod(x.shape[ax + 1:])
        oshape = x.shape[:ax] + x.shape[ax + 1:]
    out = CudaArray.empty(oshape, _I64)
    lib.mc_argmax(x.ptr, out.ptr, outer, red, inner)
    return out


def expand_reduced(g, xshape, axis, keepdims, scale=1.0):
    """Broadcast a reduction result/gradient back to `xshape` (VJP of sum)."""
    lib = L()
    axes = _norm_axes(axis, len(xshape))
    out = CudaArray.empty(xshape)
    if out.size == 0:
        return out
    if is_scalar(g):
        lib.mc_fill(out.ptr, float(g) * scale, out.size)
        return out
    g = as_f32(g)
    runs = _runs(list(axes)) if axes else []
    if len(runs) == 1:
        run = runs[0]
        outer, red, inner = _prod(xshape[:run[0]]), _prod(xshape[run[0]:run[-1] + 1]), _prod(xshape[run[-1] + 1:])
        lib.mc_bcast_mid(g.ptr, out.ptr, float(scale), outer, red, inner)
        return out
    kd = tuple(1 if i in axes else s for i, s in enumerate(xshape))
    res = binary(lib.MC_ADD, zeros(xshape), g.reshape(kd))
    return res if scale == 1.0 else binary(lib.MC_MUL, res, scale)


def select_grad(x, ans, g, axis, keepdims):
    """VJP of max/min: g broadcast * (x == ans broadcast); ties all receive g."""
    lib = L()
    axes = _norm_axes(axis, x.ndim)
    runs = _runs(list(axes)) if axes else []
    if len(runs) == 1 and not is_scalar(g):
        run = runs[0]
        outer, red, inner = _prod(x.shape[:run[0]]), _prod(x.shape[run[0]:run[-1] + 1]), _prod(x.shape[run[-1] + 1:])
        out = CudaArray.empty(x.shape)
        lib.mc_select_grad(x.ptr, as_f32(ans).ptr, as_f32(g).ptr, out.ptr, outer, red, inner)
        return out
    ansb = expand_reduced(ans, x.shape, axis, keepdims)
    gb = expand_reduced(g, x.shape, axis, keepdims)
    return binary(lib.MC_MUL, gb, binary(lib.MC_EQ, x, ansb))


def unbroadcast(g, shape):
    """Sum `g` down to `shape` (the `_unbroadcast` of numpy_core.py:119-146)."""
    lib = L()
    if g.shape == tuple(shape):
        return g
    shape = tuple(shape)
    lead = g.ndim - len(shape)
    axes = list(range(lead)) + [lead + i for i, s in enumerate(shape) if s == 1 and g.shape[lead + i] != 1]
    if not axes:
        return g.reshape(shape)
    return reduce(lib.MC_RED_SUM, g, tuple(axes)).reshape(shape)


# ------------------------------------------------------------------ shape ops ------
def reshape(x, shape):
    return as_cuda(x).reshape(shape)


def transpose(x, axes=None, out=None):
    """np.transpose as one strided copy; `out` (a float32 CudaArray or view of the result's shape) receives it in place
    of a fresh array."""
    lib = L()
    x = as_cuda(x)
    nd = x.ndim
    if axes is None:
        axes = tuple(reversed(range(nd)))
    axes = tuple(a + nd if a < 0 else a for a in axes)
    if out is None:
        if nd <= 1 or axes == tuple(range(nd)):
            return x
        if x.dtype != _F32:
            return CudaArray.from_numpy(np.transpose(x.asnumpy(), axes))
    st = _contig_strides(x.shape)
    oshape = tuple(x.shape[a] for a in axes)
    ostr = tuple(st[a] for a in axes)
    if out is None:
        out = CudaArray.empty(oshape)
    elif tuple(out.shape) != oshape or out.dtype != _F32 or x.dtype != _F32:
        raise ValueError('transpose: out must be a float32 array of shape %r' % (oshape,))
    if out.size:
        if nd > 6:
            raise ValueError('transpose of more than 6 dims is not supported')
        arr = ctypes.c_int64 * nd
        lib.mc_strided_copy(x.ptr, out.ptr, nd, arr(*oshape), arr(*ostr), 0)
    return out


def expand_dims(x, axis):
    x = as_cuda(x)
    ax = axis + x.ndim + 1 if axis < 0 else axis
    return x.reshape(x.shape[:ax] + (1,) + x.shape[ax:])


def concat2(a, b, axis):
    lib = L()
    a, b = as_f32(a), as_f32(b)
    ax = axis + a.ndim if axis < 0 else axis
    if a.shape[:ax] != b.shape[:ax] or a.shape[ax + 1:] != b.shape[ax + 1:]:
        raise ValueError('all the input array dimensions except for the concatenation axis must match exactly')
    outer, inner = _prod(a.shape[:ax]), _prod(a.shape[ax + 1:])
    oshape = a.shape[:ax] + (a.shape[ax] + b.shape[ax],) + a.shape[ax + 1:]
    out = CudaArray.empty(oshape)
    if out.size:
        lib.mc_concat_mid(a.ptr, b.ptr, out.ptr, outer, a.shape[ax], b.shape[ax], inner)
    return out


# ------------------------------------------------------------------ indexing -------
class _BasicIndex(object):
    """A basic (slice / int / None-free) index decoded to shape + strides + offset."""
    __slots__ = ('shape', 'strides', 'offset')


def decode_basic_index(xshape, index):
    """Return _BasicIndex if `index` uses only ints / slices / Ellipsis, else None."""
    if not isinstance(index, tuple):
        index = (index,)
    if any(not isinstance(i, (int, np.integer, slice, type(Ellipsis))) for i in index):
        return None
    if sum(1 for i in index if i is Ellipsis) > 1:
        raise IndexError("an index can only have a single ellipsis ('...')")
    n_spec = sum(1 for i in index if i is not Ellipsis)
    if n_spec > len(xshape):
        raise IndexError('too many indices for array')
    full = []
    for i in index:
        if i is Ellipsis:
            full.extend([slice(None)] * (len(xshape) - n_spec))
        else:
            full.append(i)
    full.extend([slice(None)] * (len(xshape) - len(full)))
    st = _contig_strides(xshape)
    r = _BasicIndex()
    shape, strides, off = [], [], 0
    for dim, (i, n) in enumerate(zip(full, xshape)):
        if isinstance(i, slice):
            start, stop, step = i.indices(n)
            cnt = max(0, (stop - start + (step - (1 if step > 0 else -1))) // step)
            shape.append(cnt)
            strides.append(st[dim] * step)
            off += start * st[dim]
        else:
            j = int(i)
            if j < -n or j >= n:
                raise IndexError('index %d is out of bounds for axis %d with size %d' % (j, dim, n))
            off += (j + n if j < 0 else j) * st[dim]
    r.shape, r.strides, r.offset = tuple(shape), tuple(strides), off
    return r


def getitem(x, index):
    """`_minpy_getitem`: basic slices on device; arange/label pairs and row gathers on device;
    anything more exotic is not part of the hot path and raises."""
    lib = L()
    x = as_cuda(x)
    bi = decode_basic_index(x.shape, index)
    if bi is not None:
        if x.dtype != _F32:
            return CudaArray.from_numpy(x.asnumpy()[index])
        out = CudaArray.empty(bi.shape)
        if out.size:
            if bi.strides == _contig_strides(bi.shape):  # contiguous block: plain copy
                lib.mc_memcpy_d2d(out.ptr, x.ptr + 4 * bi.offset, out.nbytes)
            else:
                nd = max(1, len(bi.shape))
                arr = ctypes.c_int64 * nd
                lib.mc_strided_copy(x.ptr, out.ptr, len(bi.shape), arr(*(bi.shape or (1,))), arr(*(bi.strides or (0,))),
                                    bi.offset)
        return out
    kind, payload = classify_fancy_index(x.shape, index)
    if kind == 'pick':
        idx = payload
        out = CudaArray.empty((x.shape[0],))
        lib.mc_pick_cols(as_f32(x).ptr, idx.ptr, out.ptr, x.shape[0], x.shape[1])
        return out
    if kind == 'rows':
        idx = payload
        row = _prod(x.shape[1:])
        out = CudaArray.empty(idx.shape + x.shape[1:])
        if out.size:
            lib.mc_take_rows(as_f32(x).ptr, idx.ptr, out.ptr, idx.size, row)
        return out
    raise NotImplementedError('cuda variant: unsupported index %r' % (index,))


def _is_arange(a, n):
    if isinstance(a, CudaArray):
        if a.shape != (n,):
            return False
        if a._tag is not None:
            return a._tag == ('arange', n)
        a = a.asnumpy()
    a = np.asarray(a)
    return a.shape == (n,) and a.dtype.kind in 'iuf' and np.array_equal(a, np.arange(n))


def _index_array(a, n):
    """Device int32 index array with negatives normalised."""
    if isinstance(a, CudaArray):
        if a.dtype == _I32:
            return a  # device-resident labels are trusted to be in range
        h = a.asnumpy()
    else:
        h = np.asarray(a)
    if h.dtype.kind == 'b':
        raise NotImplementedError('boolean mask indexing is not supported by the cuda variant')
    h = h.astype(np.int64)
    if h.size and (h.min() < -n or h.max() >= n):
        raise IndexError('index out of bounds for axis with size %d' % n)
    h = np.where(h < 0, h + n, h).astype(np.int32)
    return CudaArray.from_numpy(h)


def classify_fancy_index(xshape, index):
    if isinstance(index, tuple) and len(index) == 2 and len(xshape) == 2:
        r, c = index
        if not isinstance(r, slice) and not isinstance(c, slice) and not is_scalar(r) and not is_scalar(c):
            if _is_arange(r, xshape[0]):
                return 'pick', _index_array(c, xshape[1])
    if not isinstance(index, tuple) and not is_scalar(index) and not isinstance(index, slice):
        return 'rows', _index_array(index, xshape[0])
    return None, None


def getitem_grad(x, index, g):
    """`_minpy_getitem_grad`: zeros_like(x) with g added at `index` (np.add.at semantics)."""
    lib = L()
    out = zeros(x.shape)
    bi = decode_basic_index(x.shape, index)
    if is_scalar(g):
        g = full(bi.shape if bi is not None else (), float(g))
    g = as_f32(g)
    if bi is not None:
        if g.shape != bi.shape:
            g = broadcast_to(g, bi.shape)
        if g.size:
            nd = max(1, len(bi.shape))
            arr = ctypes.c_int64 * nd
            lib.mc_strided_add(out.ptr, g.ptr, len(bi.shape), arr(*(bi.shape or (1,))), arr(*(bi.strides or (0,))),
                               bi.offset)
        return out
    kind, idx = classify_fancy_index(x.shape, index)
    if kind == 'pick':
        lib.mc_pick_cols_grad(g.ptr, idx.ptr, out.ptr, x.shape[0], x.shape[1])
        return out
    if kind == 'rows':
        lib.mc_take_rows_grad(g.ptr, idx.ptr, out.ptr, idx.size, _prod(x.shape[1:]))
        return out
    raise NotImplementedError('cuda variant: unsupported index %r' % (index,))


# ------------------------------------------------------------------ dot ------------
_gemm_precision = {'mode': 1}   # 1 = TF32 tcgen05 (default), 0 = FP32 CUDA cores


def set_gemm_precision(mode):
    """'tf32' (tensor cores, default) or 'fp32' (CUDA-core exact accumulate)."""
    _gemm_precision['mode'] = {'tf32': 1, 'fp32': 0}[mode]


def gemm(a, b, trans_a=False, trans_b=False, bias=None, relu=False, out=None, accumulate=False):
    """op(a) @ op(b) for 2-D operands, optional fused bias / relu / accumulate (relu: False / True, or 2 =
    the signed-zero mode of mc_gemm)."""
    lib = L()
    a, b = as_f32(a), as_f32(b)
    if a.ndim != 2 or b.ndim != 2:
        raise ValueError('gemm expects 2-D operands')
    M, K = (a.shape[1], a.shape[0]) if trans_a else a.shape
    K2, N = (b.shape[1], b.shape[0]) if trans_b else b.shape
    if K != K2:
        raise ValueError('shapes %s and %s not aligned: %d (dim 1) != %d (dim 0)' % (a.shape, b.shape, K, K2))
    if out is None:
        out = CudaArray.empty((M, N))
    lib.mc_gemm(1 if trans_a else 0, 1 if trans_b else 0, M, N, K, a.ptr, a.shape[1], b.ptr, b.shape[1],
                out.ptr, N, bias.ptr if bias is not None else None, int(relu), 1 if accumulate else 0,
                _gemm_precision['mode'])
    return out


def dot(a, b):
    """np.dot for the cases MinPy's layers use: scalar, 1-D.1-D, 2-D.1-D, 1-D.2-D, 2-D.2-D, N-D.2-D."""
    lib = L()
    if is_scalar(a) or is_scalar(b):
        return binary(lib.MC_MUL, a, b)
    a, b = as_f32(a), as_f32(b)
    if a.ndim == 2 and b.ndim == 2:
        return gemm(a, b)
    if a.ndim == 1 and b.ndim == 1:
        return reduce(lib.MC_RED_SUM, binary(lib.MC_MUL, a, b))
    if a.ndim == 1 and b.ndim == 2:
        return gemm(a.reshape(1, a.shape[0]), b).reshape(b.shape[1])
    if b.ndim == 1:
        lead = a.shape[:-1]
        return gemm(a.reshape(_prod(lead), a.shape[-1]), b.reshape(b.shape[0], 1)).reshape(lead)
    if b.ndim == 2:
        lead = a.shape[:-1]
        return gemm(a.reshape(_prod(lead), a.shape[-1]), b).reshape(lead + (b.shape[1],))
    raise NotImplementedError('dot of %d-D with %d-D arrays is not supported by the cuda variant' % (a.ndim, b.ndim))


# ------------------------------------------------------------------ softmax --------
def softmax_rows(x):
    x = as_f32(x)
    rows, cols = _prod(x.shape[:-1]), x.shape[-1]
    out = CudaArray.empty(x.shape)
    if out.size:
        L().mc_softmax_rows(x.ptr, out.ptr, rows, cols)
    return out


def logsumexp_rows(x):
    x = as_f32(x)
    rows, cols = _prod(x.shape[:-1]), x.shape[-1]
    out = CudaArray.empty(x.shape[:-1])
    L().mc_logsumexp_rows(x.ptr, out.ptr, rows, cols)
    return out


def onehot_encode(indices, out):
    """In place: out[i, indices[i]] = 1 (numpy_core.py:47-63)."""
    idx = as_i32(indices)
    L().mc_onehot(idx.ptr, out.ptr, out.shape[0], out.shape[1])
    return out


# ------------------------------------------------------------------ VJP chain ------
def run_chain(g, g_scalar, steps, shape, acc=None, g_mode=None, g_inner=1):
    """out = (acc or 0) + g * g_scalar * prod_k f_k, elementwise over `shape`, in ONE kernel
    (one per 12 steps for longer recipes).

    `g` is a CudaArray addressed by `g_mode` (FULL / ROW / COL / ONE; default FULL, or ONE
    when g has a single element) or None (then `g_scalar` alone seeds the chain); `steps` is
    a list of tuples (op, a, a_mode, b, b_mode, scalar, inner)."""
    lib = L()
    out = CudaArray.empty(shape)
    n = out.size
    if n == 0:
        return out
    if g is not None and g_mode is None:
        g_mode = 0 if g.size == n else 3
    nchunks = max(1, (len(steps) + 11) // 12)
    for c in range(nchunks):
        chunk = steps[c * 12:(c + 1) * 12]
        arr = (VjpStep * max(1, len(chunk)))()
        for i, (op, a, am, b, bm, s, inner) in enumerate(chunk):
            arr[i].op, arr[i].a_mode, arr[i].b_mode, arr[i].scalar = op, am, bm, s
            arr[i].a = a.ptr if a is not None else None
            arr[i].b = b.ptr if b is not None else None
            arr[i].inner = inner
        last = c == nchunks - 1
        lib.mc_vjp_chain(g.ptr if g is not None else None, g_mode or 0, g_inner, float(g_scalar),
                         acc.ptr if (acc is not None and last) else None, arr, len(chunk), out.ptr, n)
        if not last:
            g, g_scalar, g_mode, g_inner = out, 1.0, 0, 1
            out = CudaArray.empty(shape)
    return out


# ------------------------------------------------------------------ convolution / pooling ---
def conv2d_fwd(x, w, b=None, stride=1, pad=0, relu=False):
    """NCHW cross-correlation y = conv(x, w) + b [relu]: the `Convolution` (+ `Activation`)
    symbols of examples/nn/cnn.py:33-36."""
    lib = L()
    x, w = as_f32(x), as_f32(w)
    if x.ndim != 4 or w.ndim != 4 or x.shape[1] != w.shape[1]:
        raise ValueError('conv2d: x %s and w %s do not match (NCHW / FCHW)' % (x.shape, w.shape))
    N, C, H, W = x.shape
    F, _, KH, KW = w.shape
    OH, OW = (H + 2 * pad - KH) // stride + 1, (W + 2 * pad - KW) // stride + 1
    if OH <= 0 or OW <= 0:
        raise ValueError('conv2d: kernel larger than padded input')
    y = CudaArray.empty((N, F, OH, OW))
    if y.size:
        lib.mc_conv2d_fwd(x.ptr, w.ptr, as_f32(b).ptr if b is not None else None, y.ptr, N, C, H, W, F, KH, KW,
                          stride, pad, 1 if relu else 0, _gemm_precision['mode'])
    return y


def conv2d_bwd(x, w, dy, stride=1, pad=0, need_dx=True, need_db=True):
    """(dx, dw, db) of conv2d_fwd; dx is None when `need_dx` is False (the data input of the
    first layer is not on the tape)."""
    lib = L()
    x, w, dy = as_f32(x), as_f32(w), as_f32(dy)
    N, C, H, W = x.shape
    F, _, KH, KW = w.shape
    dx = None
    if need_dx:
        dx = CudaArray.empty(x.shape)
        if dx.size:
            lib.mc_conv2d_dgrad(dy.ptr, w.ptr, dx.ptr, N, C, H, W, F, KH, KW, stride, pad, _gemm_precision['mode'])
    dw = CudaArray.empty(w.shape)
    db = CudaArray.empty((F,)) if need_db else None
    lib.mc_conv2d_wgrad(x.ptr, dy.ptr, dw.ptr, db.ptr if db is not None else None, N, C, H, W, F, KH, KW, stride, pad,
                        _gemm_precision['mode'])
    return dx, dw, db


def maxpool2d_fwd(x, kernel, stride, want_argmax=True):
    """'valid' max pooling over NCHW; returns (y, argmax int32 or None)."""
    lib = L()
    x = as_f32(x)
    N, C, H, W = x.shape
    KH, KW = (kernel, kernel) if isinstance(kernel, int) else kernel
    OH, OW = (H - KH) // stride + 1, (W - KW) // stride + 1
    if OH <= 0 or OW <= 0:
        raise ValueError('maxpool2d: kernel larger than input')
    y = CudaArray.empty((N, C, OH, OW))
    arg = CudaArray.empty((N, C, OH, OW), _I32) if want_argmax else None
    if y.size:
        lib.mc_maxpool2d_fwd(x.ptr, y.ptr, arg.ptr if arg is not None else None, N * C, H, W, KH, KW, stride)
    return y, arg


def maxpool2d_bwd(dy, arg, xshape):
    """Routes each output gradient to its window's arg-max element (overlapping windows add)."""
    lib = L()
    dy = as_f32(dy)
    N, C, H, W = xshape
    dx = zeros(xshape)
    if dy.size:
        lib.mc_maxpool2d_bwd(dy.ptr, arg.ptr, dx.ptr, N * C, H, W, dy.shape[2], dy.shape[3])
    return dx


def avgpool2d_fwd(x, kernel, stride):
    """'valid' average pooling over NCHW."""
    x = as_f32(x)
    N, C, H, W = x.shape
    KH, KW = (kernel, kernel) if isinstance(kernel, int) else kernel
    OH, OW = (H - KH) // stride + 1, (W - KW) // stride + 1
    if OH <= 0 or OW <= 0:
        raise ValueError('avgpool2d: kernel larger than input')
    y = CudaArray.empty((N, C, OH, OW))
    if y.size:
        L().mc_avgpool2d_fwd(x.ptr, y.ptr, N * C, H, W, KH, KW, stride)
    return y


def avgpool2d_bwd(dy, xshape, kernel, stride):
    dy = as_f32(dy)
    N, C, H, W = xshape
    KH, KW = (kernel, kernel) if isinstance(kernel, int) else kernel
    dx = CudaArray.empty(xshape)
    if dx.size:
        L().mc_avgpool2d_bwd(dy.ptr, dx.ptr, N * C, H, W, KH, KW, stride)
    return dx


# ------------------------------------------------------- channels-last (NHWC) conv trunk ----
class Nhwc(object):
    """A channels-last activation: buffer (N, H, W + 2*wpad, Cp) for the logical NCHW tensor
    (N, C, H, W).  What the tensor-core convolution consumes and produces (csrc/conv_tc.cu)."""
    __slots__ = ('data', 'N', 'C', 'H', 'W', 'Cp', 'wpad')

    def __init__(self, data, N, C, H, W, Cp, wpad=0):
        self.data, self.N, self.C, self.H, self.W, self.Cp, self.wpad = data, N, C, H, W, Cp, wpad

    @property
    def shape(self):
        return (self.N, self.C, self.H, self.W)


def _round4(c):
    return (c + 3) // 4 * 4


def to_nhwc(x, wpad=0):
    """NCHW CudaArray -> Nhwc (zero pad pixels / pad channels)."""
    x = as_f32(x)
    N, C, H, W = x.shape
    Cp = _round4(C)
    out = CudaArray.empty((N, H, W + 2 * wpad, Cp))
    if out.size:
        L().mc_nchw_to_nhwc(x.ptr, out.ptr, N, C, H, W, Cp, wpad)
    return Nhwc(out, N, C, H, W, Cp, wpad)


def to_nchw(a):
    """Nhwc (un-padded) -> NCHW CudaArray."""
    if a.wpad:
        raise ValueError('to_nchw: W-padded activations are inputs only')
    out = CudaArray.empty((a.N, a.C, a.H, a.W))
    if out.size:
        L().mc_nhwc_to_nchw(a.data.ptr, out.ptr, a.N, a.C, a.H, a.W, a.Cp)
    return out


def conv_merged_ok(C, KW):
    """Few-channel 'merged' mode (one k-block per filter row) needs KW * Cp <= 32."""
    return KW * _round4(C) <= 32


def conv2d_nhwc_fwd(a, w, b=None, pad=0, relu=False):
    lib = L()
    w = as_f32(w)
    F, C, KH, KW = w.shape
    if C != a.C:
        raise ValueError('conv2d: input has %d channels, filter expects %d' % (a.C, C))
    OH, OW = a.H + 2 * pad - KH + 1, a.W + 2 * pad - KW + 1
    if OH <= 0 or OW <= 0:
        raise ValueError('conv2d: kernel larger than padded input')
    Fp = _round4(F)
    y = CudaArray.empty((a.N, OH, OW, Fp))
    if y.size:
        lib.mc_conv2d_nhwc_fwd(a.data.ptr, w.ptr, as_f32(b).ptr if b is not None else None, y.ptr, a.N, C, a.H, a.W,
                               a.Cp, a.wpad, F, Fp, KH, KW, pad, 1 if relu else 0)
    return Nhwc(y, a.N, F, OH, OW, Fp)


def conv_pool_fusable(a, w, pad):
    """mc_conv2d_nhwc_pool_supported, restated: conv output even-sized, at most 32 wide, F a multiple of 32."""
    F, C, KH, KW = w.shape
    OH, OW = a.H + 2 * pad - KH + 1, a.W + 2 * pad - KW + 1
    return OH > 0 and OW > 0 and OH % 2 == 0 and OW % 2 == 0 and OW <= 32 and F % 32 == 0


def conv2d_nhwc_fwd_pool(a, w, b=None, pad=0, relu=True, want_idx=True):
    """conv -> (relu) -> 2x2 / 2 max pooling in ONE kernel: returns (pooled Nhwc, uint8 window index or None,
    `like` = an Nhwc describing the un-pooled convolution output, which is never stored)."""
    lib = L()
    w = as_f32(w)
    F, C, KH, KW = w.shape
    if C != a.C:
        raise ValueError('conv2d: input has %d channels, filter expects %d' % (a.C, C))
    OH, OW = a.H + 2 * pad - KH + 1, a.W + 2 * pad - KW + 1
    y = CudaArray.empty((a.N, OH // 2, OW // 2, F))
    idx = CudaArray.empty((a.N, OH // 2, OW // 2, F), np.dtype('uint8')) if want_idx else None
    if y.size:
        lib.mc_conv2d_nhwc_fwd_pool(a.data.ptr, w.ptr, as_f32(b).ptr if b is not None else None, y.ptr,
                                    idx.ptr if idx is not None else None, a.N, C, a.H, a.W, a.Cp, a.wpad, F, F, KH, KW, pad,
                                    1 if relu else 0)
    return Nhwc(y, a.N, F, OH // 2, OW // 2, F), idx, Nhwc(None, a.N, F, OH, OW, F)


def conv2d_nhwc_bwd(a, w, dy, pad=0, need_dx=True, need_db=True):
    """(dx Nhwc or None, dw, db) for y = conv(a, w); dy is the Nhwc gradient of y."""
    lib = L()
    w = as_f32(w)
    F, C, KH, KW = w.shape
    dx = None
    if need_dx:
        d = CudaArray.empty((a.N, a.H, a.W, a.Cp))
        if d.size:
            lib.mc_conv2d_nhwc_dgrad(dy.data.ptr, w.ptr, d.ptr, a.N, C, a.H, a.W, a.Cp, F, dy.Cp, KH, KW, pad)
        dx = Nhwc(d, a.N, C, a.H, a.W, a.Cp)
    dw = CudaArray.empty(w.shape)
    db = CudaArray.empty((F,)) if need_db else None
    # the weight gradient is read only by the update rule: beside the data-gradient chain, on the side stream
    with side_stream(a.data, dy.data, dw, db, any_size=True):     # measured: cnn_c4 0.2937 beside the chain, 0.295 in sequence
        lib.mc_conv2d_nhwc_wgrad(a.data.ptr, dy.data.ptr, dw.ptr, db.ptr if db is not None else None, a.N, C, a.H, a.W,
                                 a.Cp, a.wpad, F, dy.Cp, KH, KW, pad)
    return dx, dw, db


def maxpool2d_nhwc_fwd(a, k, want_idx=True):
    """k x k / stride k max pooling on an Nhwc activation; returns (Nhwc, uint8 window index)."""
    OH, OW = (a.H - k) // k + 1, (a.W - k) // k + 1
    if OH <= 0 or OW <= 0:
        raise ValueError('maxpool2d: kernel larger than input')
    y = CudaArray.empty((a.N, OH, OW, a.Cp))
    idx = CudaArray.empty((a.N, OH, OW, a.Cp), np.dtype('uint8')) if want_idx else None
    if y.size:
        L().mc_maxpool2d_nhwc_fwd(a.data.ptr, y.ptr, idx.ptr if idx is not None else None, a.N, a.H, a.W, a.Cp, k, k, k)
    return Nhwc(y, a.N, a.C, OH, OW, a.Cp), idx


def maxpool2d_nhwc_bwd(dy, idx, like, k, relu_y=None):
    """Gradient of maxpool2d_nhwc_fwd w.r.t. its input `like`; `relu_y` (the pooled output) folds
    the producing relu's mask into the same pass."""
    dx = CudaArray.empty((like.N, like.H, like.W, like.Cp))
    if dx.size:
        L().mc_maxpool2d_nhwc_bwd(dy.data.ptr, idx.ptr, relu_y.data.ptr if relu_y is not None else None, dx.ptr,
                                  like.N, like.H, like.W, like.Cp, k, dy.H, dy.W)
    return Nhwc(dx, like.N, like.C, like.H, like.W, like.Cp)


# ------------------------------------------------------------------ side stream ----
import contextlib as _contextlib
import os as _os

SIDE_STREAM = _os.environ.get('MINPY_SIDE_STREAM', '1') != '0'
SIDE_MAX_BYTES = 32 << 20      # operands of one side-stream section beyond this run on the compute stream
_side = {'keep': [], 'pending': False, 'active': False}


@_contextlib.contextmanager
def side_stream(*operands, any_size=False):
    """Library calls inside the block run on the side stream (mc_side_begin / mc_side_end): for work whose
    result nothing reads before the end of the backward pass (weight gradients).  `operands` -- everything the
    block reads -- and whatever is passed to `side_keep` stay referenced until `side_join`, because the
    host-side free lists hand a dropped buffer to the next MAIN-stream kernel at once."""
    # only for work that leaves room beside it: a contraction over tens of MB fills every SM and the memory system by
    # itself (lstm_c5: its weight-gradient GEMMs beside the data-gradient chain measured 1.7 % SLOWER than in sequence)
    big = (not any_size) and sum(getattr(a, 'nbytes', 0) for a in operands) > SIDE_MAX_BYTES
    if not SIDE_STREAM or _side['active'] or big:
        yield False
        return
    lib = L()
    _side['keep'].extend(operands)
    lib.mc_side_begin()
    _side['active'] = True
    try:
        yield True
    finally:
        _side['active'] = False
        _side['pending'] = True
        lib.mc_side_end()


def side_keep(*arrays):
    if _side['pending'] or _side['active']:
        _side['keep'].extend(arrays)


def side_join():
    """The compute stream waits for all side work issued so far; its operands may be released."""
    if _side['pending']:
        L().mc_side_join()
        _side['pending'] = False
        del _side['keep'][:]


# ------------------------------------------------------- operators of the raw array type ----
def _install_operators():
    """The raw-array duck type MinPy's core and user-written VJPs / `customop('cuda')` bodies use
    (arithmetic operators, unary minus, comparisons as methods, `.sum() .max() .T`,
    `.broadcast_to`; indexing is `ops.getitem` -- a `__getitem__` would make NumPy treat the object as a sequence).  `==` stays identity (CudaArrays live in dicts / sets of the step
    graph); elementwise equality is `ops.binary(L().MC_EQ, a, b)`."""
    def bop(name):
        def fwd(a, b):
            return binary(getattr(L(), name), a, b)

        def rev(a, b):
            return binary(getattr(L(), name), b, a)
        return fwd, rev
    for py, mc in (('add', 'MC_ADD'), ('sub', 'MC_SUB'), ('mul', 'MC_MUL'), ('truediv', 'MC_DIV'), ('pow', 'MC_POW'),
                   ('mod', 'MC_MOD')):
        fwd, rev = bop(mc)
        setattr(CudaArray, '__%s__' % py, fwd)
        setattr(CudaArray, '__r%s__' % py, rev)
    for py, mc in (('lt', 'MC_LT'), ('gt', 'MC_GT'), ('le', 'MC_LE'), ('ge', 'MC_GE')):
        setattr(CudaArray, '__%s__' % py, bop(mc)[0])
    CudaArray.__neg__ = lambda a: unary(L().MC_NEG, a)
    CudaArray.__abs__ = lambda a: unary(L().MC_ABS, a)
    CudaArray.sum = lambda a, axis=None, keepdims=False: reduce(L().MC_RED_SUM, a, axis, keepdims)
    CudaArray.max = lambda a, axis=None, keepdims=False: reduce(L().MC_RED_MAX, a, axis, keepdims)
    CudaArray.min = lambda a, axis=None, keepdims=False: reduce(L().MC_RED_MIN, a, axis, keepdims)
    CudaArray.transpose = lambda a, *axes: transpose(a, axes[0] if len(axes) == 1 and not is_scalar(axes[0]) else (axes or None))
    CudaArray.T = property(lambda a: transpose(a))
    CudaArray.broadcast_to = lambda a, shape: broadcast_to(a, shape)
    CudaArray.dot = lambda a, b: dot(a, b)


_install_operators()
