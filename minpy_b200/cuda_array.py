"""CudaArray: the raw array type of the `cuda` array variant.

Plays the role `mxnet.ndarray.NDArray` plays for the reference's MXNet variant
(minpy/array_variants/mxnet/__init__.py:10): a dense, row-major device buffer with a
shape and a dtype, plus the duck-typed surface MinPy's core calls on a raw array
(`.shape .dtype .size .asnumpy() .wait_to_read() .reshape() .copyto()`,
SURVEY.md 8b).  Storage comes from the library's stream-ordered pool; it is released
when the last view dies.
"""
from __future__ import annotations

import ctypes

import numpy as np

from ._lib import get_lib

_F32 = np.dtype('float32')
_I32 = np.dtype('int32')
_I64 = np.dtype('int64')
_DTYPE_CODE = {_F32: 0, _I32: 1, _I64: 2, np.dtype('float64'): 3, np.dtype('uint8'): 4, np.dtype('bool'): 4}

_state = {'device': None, 'home': None}


def ensure_device(device: int = None):
    """Make `device` the library's current device (default: the process's home device = LOCAL_RANK
    or 0, chosen at the first call).  Switching is cheap: each device keeps its own stream."""
    cur = _state['device']
    if cur is None or (device is not None and device != cur):
        import os
        dev = device if device is not None else int(os.environ.get('LOCAL_RANK', '0'))
        get_lib().mc_init(dev)
        _state['device'] = dev
        if _state['home'] is None:
            _state['home'] = dev
        return dev
    return cur


def enter_context_device(ctx):
    """Select the device a MinPy context names (`with gpu(i):`, minpy/context.py:92-98); a cpu context
    means "no placement request": the process's home device."""
    want = ctx.cuda_device() if ctx is not None else None
    if want is None:
        want = _state['home']
        if want is None:
            return ensure_device()
    if want != _state['device']:
        ensure_device(want)
    return want


class BufferPool(object):
    """Blocks reserved for the temporaries of one captured step graph."""
    __slots__ = ('lists', 'released')

    def __init__(self):
        self.lists = {}
        self.released = False

    def release(self):
        """The graph is gone: the blocks rejoin the general free list."""
        self.released = True
        lists, self.lists = self.lists, {}
        for key, blocks in lists.items():
            nbytes = key[1] if isinstance(key, tuple) else key
            _Buffer._free_lists.setdefault(key, []).extend(blocks)      # pylint: disable=protected-access
            _Buffer._cached_bytes += nbytes * len(blocks)               # pylint: disable=protected-access


class _Buffer(object):
    """Owns one device allocation.

    Released blocks go to a host-side free list keyed by (rounded) size and are handed out again
    without calling into the library: every kernel of this process is enqueued on ONE stream, so
    reuse in program order is reuse in stream order -- the same guarantee the library's
    cudaMallocAsync pool gives, minus two ctypes round trips per temporary (a training step
    creates ~100 temporaries of a handful of distinct sizes).  The list is bounded; beyond the
    bound blocks really go back to the library."""
    __slots__ = ('ptr', 'nbytes', 'pool', 'device')
    _free_lists = {}
    _cached_bytes = 0
    _cache_limit = 32 << 30
    _lib_alloc = None
    _lib_free = None
    _active_pool = None      # a BufferPool while a step graph is being captured (step_graph.py)

    def __init__(self, nbytes):
        nbytes = (nbytes + 511) & ~511 if nbytes else 512
        self.nbytes = nbytes
        dev = self.device = _state['device']
        # blocks of a `with gpu(i):` scope off the home device live in free lists of their own
        key = nbytes if dev == _state['home'] else (dev, nbytes)
        pool = self.pool = _Buffer._active_pool
        if pool is not None:
            free = pool.lists.get(key)
            if free:
                self.ptr = free.pop()
                return
        else:
            free = _Buffer._free_lists.get(key)
            if free:
                self.ptr = free.pop()
                _Buffer._cached_bytes -= nbytes
                return
        if _Buffer._lib_alloc is None:
            lib = get_lib()
            _Buffer._lib_alloc, _Buffer._lib_free = lib.mc_alloc, lib.cdll.mc_free
        p = ctypes.c_void_p()
        try:
            _Buffer._lib_alloc(ctypes.byref(p), nbytes)
        except Exception:
            _Buffer.trim()                  # give the cached blocks back and retry once
            _Buffer._lib_alloc(ctypes.byref(p), nbytes)
        self.ptr = p.value

    def __del__(self):
        try:
            pool = self.pool
            key = self.nbytes if self.device == _state['home'] else (self.device, self.nbytes)
            if pool is not None and not pool.released:
                # a block a captured graph writes on every replay: it may only ever be handed to
                # another capture-time temporary of the same graph, whenever this object dies
                free = pool.lists.get(key)
                if free is None:
                    free = pool.lists[key] = []
                free.append(self.ptr)
            elif _Buffer._cached_bytes + self.nbytes <= _Buffer._cache_limit or self.device != _state['device']:
                # (a block of another device is always cached: freeing needs that device's stream)
                free = _Buffer._free_lists.get(key)
                if free is None:
                    free = _Buffer._free_lists[key] = []
                free.append(self.ptr)
                _Buffer._cached_bytes += self.nbytes
            else:
                _Buffer._lib_free(self.ptr)
        except Exception:  # interpreter shutdown
            pass

    @staticmethod
    def trim():
        """Return every cached block to the library's pool (each on the device that owns it)."""
        lists, _Buffer._free_lists = _Buffer._free_lists, {}
        _Buffer._cached_bytes = 0
        cur = _state['device']
        for key, blocks in lists.items():
            dev = key[0] if isinstance(key, tuple) else _state['home']
            if dev != _state['device']:
                ensure_device(dev)
            for ptr in blocks:
                _Buffer._lib_free(ptr)
        if cur is not None and cur != _state['device']:
            ensure_device(cur)


def _prod(shape):
    n = 1
    for s in shape:
        n *= s
    return n


class CudaArray(object):
    """Dense row-major device array (float32 / int32 / int64)."""
    __slots__ = ('buf', 'ptr', 'shape', 'size', 'dtype', '_saved', '_tag', '_readback', '__weakref__')
    __array_priority__ = 200.0

    def __init__(self, buf, ptr, shape, dtype=_F32):
        self.buf = buf
        self.ptr = ptr
        self.shape = shape
        self.size = _prod(shape)
        self.dtype = dtype
        self._saved = None    # forward state a fused primitive keeps for its VJP
        self._tag = None      # ('arange', n) for index vectors known to be 0..n-1
        self._readback = None  # ticket of a device->host copy already in flight (begin_readback)

    # ---- construction -------------------------------------------------------------
    @staticmethod
    def empty(shape, dtype=_F32):
        if _state['device'] is None:
            ensure_device()
        shape = tuple(int(s) for s in shape)
        buf = _Buffer(_prod(shape) * dtype.itemsize)
        return CudaArray(buf, buf.ptr, shape, dtype)

    @staticmethod
    def from_numpy(arr):
        """Host -> device. float64/float16 become float32 (the MXNet variant did the same,
        minpy/array.py:388-389); integer and bool arrays become int32/int64-preserving."""
        arr = np.asarray(arr)
        if arr.dtype.kind == 'f':
            dt = _F32
        elif arr.dtype.kind == 'b':
            dt = _F32
        elif arr.dtype == _I64:
            dt = _I64
        elif arr.dtype.kind in 'iu':
            dt = _I32
        else:
            raise TypeError('unsupported dtype for the cuda variant: %s' % arr.dtype)
        host = np.ascontiguousarray(arr, dtype=dt)
        out = CudaArray.empty(host.shape, dt)
        if host.size:
            get_lib().mc_memcpy_h2d(out.ptr, host.ctypes.data, host.nbytes)
            # pageable source: the copy is staged before the call returns, `host` may die
        return out

    # ---- duck-typed surface ---------------------------------------------------------
    @property
    def ndim(self):
        return len(self.shape)

    @property
    def nbytes(self):
        return self.size * self.dtype.itemsize

    @property
    def device(self):
        """CUDA ordinal of the device that holds the data."""
        return self.buf.device

    @property
    def context(self):
        return ('gpu', self.buf.device)

    def to_device(self, device):
        """A copy on `device` of this process (the current device is switched to it)."""
        device = int(device)
        ensure_device(device)
        if self.buf.device == device:
            return self
        out = CudaArray.empty(self.shape, self.dtype)
        get_lib().mc_memcpy_peer(out.ptr, device, self.ptr, self.buf.device, self.nbytes)
        return out

    def begin_readback(self):
        """Start copying a small value (<= 64 bytes, never written again) to the host on the
        library's readback stream; a later asnumpy() / float() waits for that copy only, not for
        compute work enqueued after this call (the Solver does this for every step's loss)."""
        if 0 < self.nbytes <= 64:
            ticket = ctypes.c_uint64()
            get_lib().mc_readback_begin(self.ptr, self.nbytes, ctypes.byref(ticket))
            self._readback = ticket.value
        return self

    def asnumpy(self):
        if self.buf.device != _state['device']:
            ensure_device(self.buf.device)        # the copy runs on the owning device's stream
        out = np.empty(self.shape, dtype=self.dtype)
        if self._readback is not None:
            lib = get_lib()
            if lib.mc_readback_wait.raw(self._readback, out.ctypes.data, self.nbytes) == 0:
                return out
            self._readback = None     # the ring moved on (status 4): blocking read below
        get_lib().mc_memcpy_d2h(out.ctypes.data, self.ptr, self.nbytes)
        return out

    def asscalar(self):
        return self.asnumpy().reshape(-1)[0].item()

    def wait_to_read(self):
        get_lib().mc_sync()

    def reshape(self, *shape):
        if len(shape) == 1 and not isinstance(shape[0], (int, np.integer)):
            shape = tuple(shape[0])
        shape = [int(s) for s in shape]
        if -1 in shape:
            known = -_prod(shape)
            shape[shape.index(-1)] = self.size // known if known else 0
        shape = tuple(shape)
        if _prod(shape) != self.size:
            raise ValueError('cannot reshape array of size %d into shape %s' % (self.size, shape))
        return CudaArray(self.buf, self.ptr, shape, self.dtype)

    def copy(self):
        out = CudaArray.empty(self.shape, self.dtype)
        get_lib().mc_memcpy_d2d(out.ptr, self.ptr, self.nbytes)
        return out

    def copyto(self, dst):
        if dst.shape != self.shape or dst.dtype != self.dtype:
            raise ValueError('copyto: shape/dtype mismatch')
        get_lib().mc_memcpy_d2d(dst.ptr, self.ptr, self.nbytes)
        return dst

    def astype(self, dtype):
        dtype = np.dtype(dtype)
        if dtype == np.dtype('float64'):
            dtype = _F32
        if dtype == self.dtype:
            return self
        out = CudaArray.empty(self.shape, dtype)
        get_lib().mc_cast(self.ptr, _DTYPE_CODE[self.dtype], out.ptr, _DTYPE_CODE[dtype], self.size)
        return out

    def __len__(self):
        if not self.shape:
            raise TypeError('len() of unsized object')
        return self.shape[0]

    def __repr__(self):
        return 'CudaArray(shape=%s, dtype=%s)' % (self.shape, self.dtype)

    def __float__(self):
        return float(self.asnumpy().reshape(-1)[0])

    # operators are provided by minpy_b200.ops (installed there to avoid an import cycle)
