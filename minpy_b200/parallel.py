"""Data parallelism for `Solver._step`: one process per GPU, gradients averaged across ranks.

The reference has no multi-device training loop at all (only manual
`with gpu(i)` placement, minpy/context.py:92-98).  Minibatch training shards by rows and every
loss in nn/layers.py is a mean over the batch, so the full-batch gradient is the average of
the per-shard gradients: ONE exchange step per training step, placed between the tape's
backward pass and the update rules (minpy/nn/solver.py:184 and :190 in the reference).

Plumbing is `torch.distributed` (NCCL over NVLink on the device, gloo on the host for the
CPU tests).  The gradients are packed into one persistent flat device buffer with the
library's multi-tensor pack kernel, all-reduced in place (a 0.5-20 MB message: latency bound,
so one message, not one per parameter), and handed back as zero-copy views of the flat buffer.
The library and torch share the legacy default stream, so ordering needs no extra events.
"""
from __future__ import annotations

import ctypes
import os
import sys

import numpy as np

_state = {'initialised': False, 'world': 1, 'rank': 0, 'backend': None, 'flat': None, 'flat_t': None, 'native': False}


def _dist():
    import torch.distributed as dist
    return dist


def init(backend=None):
    """Join the process group described by RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT
    (torchrun sets them).  A single process (no WORLD_SIZE, or WORLD_SIZE=1) is a no-op."""
    if _state['initialised']:
        return _state['world']
    world = int(os.environ.get('WORLD_SIZE', '1'))
    if world <= 1:
        _state.update(initialised=True, world=1, rank=0)
        return 1
    import torch
    dist = _dist()
    if backend is None:
        backend = 'nccl' if torch.cuda.is_available() else 'gloo'
    os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
    os.environ.setdefault('MASTER_PORT', '29500')
    rank = int(os.environ.get('RANK', '0'))
    if backend == 'nccl':
        local = int(os.environ.get('LOCAL_RANK', str(rank)))
        torch.cuda.set_device(local)
        from .cuda_array import ensure_device
        ensure_device(local)
    if not dist.is_initialized():
        kwargs = {}
        if backend == 'nccl':
            kwargs['device_id'] = torch.device('cuda', int(os.environ.get('LOCAL_RANK', str(rank))))
        dist.init_process_group(backend=backend, rank=rank, world_size=world, **kwargs)
    _state.update(initialised=True, world=world, rank=rank, backend=backend)
    if backend == 'nccl' and os.environ.get('MINPY_NATIVE_NCCL', '1') == '1':
        _init_native_comm(rank, world)
    return world


def _init_native_comm(rank, world):
    """The library's own NCCL communicator (csrc/comm.cu): collectives on the library stream, hence
    inside the captured step graph.  torch.distributed stays the rendezvous (it carries the 128-byte
    unique id) and the barrier."""
    from .ops import L
    lib = L()
    dist = _dist()
    uid = ctypes.create_string_buffer(128)
    if rank == 0:
        lib.mc_comm_unique_id(uid)
    box = [uid.raw]
    dist.broadcast_object_list(box, src=0)
    lib.mc_comm_init(rank, world, ctypes.create_string_buffer(box[0], 128))
    _state['native'] = True
    _init_p2p_window(rank, world)


def _init_p2p_window(rank, world):
    """Peer-memory window for the gradient bucket (csrc/comm.cu, two-shot all-reduce over NVLink peer
    loads): every rank allocates it, the IPC handles travel through torch.distributed, and the window
    is used only if EVERY rank could map every peer.  MINPY_P2P_ALLREDUCE: 'auto' (default) and '1' use the
    window at every world size 2..8 -- with the push kernel (round 2) it beats NCCL alone and inside the step
    (9.5 MB: 32 vs 48 us at 2 ranks, 44 vs 52 us at 4; mlp_c3 step 0.2195 vs 0.2467 ms at 2 ranks, 0.2303 vs
    0.2518 at 4: profiles/r2n2b, r2n4; the round-1 pull kernel only paid off at 2 ranks); '0' keeps NCCL.
    MINPY_P2P_MB sizes the window (default 64 MB, buckets beyond it go to NCCL)."""
    _state['p2p'] = None
    mode = os.environ.get('MINPY_P2P_ALLREDUCE', 'auto')
    if world < 2 or world > 8 or mode == '0':
        return
    from .ops import L
    lib = L()
    dist = _dist()
    nbytes = int(float(os.environ.get('MINPY_P2P_MB', '64')) * (1 << 20))
    handles = ctypes.create_string_buffer(128)
    ok, err = True, ''
    try:
        lib.mc_p2p_create(rank, world, nbytes, handles)
    except Exception as exc:  # pylint: disable=broad-except
        ok, err = False, str(exc)
    gathered = [None] * world
    dist.all_gather_object(gathered, (ok, handles.raw))
    if not all(g[0] for g in gathered):
        if rank == 0:
            print('minpy_b200.parallel: peer-memory window unavailable (%s); gradients go through NCCL' % err,
                  file=sys.stderr)
        return
    try:
        lib.mc_p2p_connect(ctypes.create_string_buffer(b''.join(g[1] for g in gathered), 128 * world))
    except Exception as exc:  # pylint: disable=broad-except
        ok, err = False, str(exc)
    flags = [None] * world
    dist.all_gather_object(flags, ok)
    if not all(flags):
        # a rank that did connect must not use the window alone (the kernel's barriers need everyone):
        # _state['p2p'] stays None on every rank, so no bucket is ever placed in the window
        if rank == 0:
            print('minpy_b200.parallel: peer mapping failed on a rank (%s); gradients go through NCCL' % err,
                  file=sys.stderr)
        return
    ptr, size, bad = ctypes.c_void_p(), ctypes.c_size_t(), ctypes.c_int()
    lib.mc_p2p_info(ctypes.byref(ptr), ctypes.byref(size), ctypes.byref(bad))
    if ptr.value:
        _state['p2p'] = (ptr.value, size.value)


class _WindowBuffer(object):
    """Stand-in for cuda_array._Buffer over the peer-memory window: owned by the library, never freed."""
    __slots__ = ('ptr', 'nbytes', 'pool', 'device')

    def __init__(self, ptr, nbytes):
        from .cuda_array import ensure_device
        self.ptr, self.nbytes, self.pool, self.device = ptr, nbytes, None, ensure_device()


def p2p_active():
    """True when the gradient bucket is exchanged by the library's own peer-memory all-reduce."""
    return _state.get('p2p') is not None


def native_comm():
    """True when gradients are exchanged by the library's own communicator (capturable)."""
    return bool(_state['native'])


def shutdown():
    # The library communicator is NOT destroyed here: step graphs that captured its all-reduce keep
    # it referenced and ncclCommDestroy then blocks forever (seen on 2 x B200, gpurun_out/r1n); the
    # process exit reclaims it.  mc_comm_destroy stays available for callers that own no graphs.
    if _state['world'] > 1 and _dist().is_initialized():
        _dist().destroy_process_group()
    _state.update(initialised=False, world=1, rank=0, backend=None, flat=None, flat_t=None, native=False, p2p=None)


def world_size():
    return _state['world']


def rank():
    return _state['rank']


def barrier():
    if _state['world'] > 1:
        _dist().barrier()


class _DevMem(object):
    """`__cuda_array_interface__` view of a device allocation, so torch can alias it."""

    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {'shape': (n,), 'typestr': '<f4', 'data': (ptr, False), 'version': 2}


def torch_view(arr):
    """torch.Tensor aliasing a float32 CudaArray (no copy)."""
    import torch
    from .cuda_array import ensure_device
    return torch.as_tensor(_DevMem(arr.ptr, max(1, arr.size)), device='cuda:%d' % ensure_device())[:arr.size]


_ALIGN = 64   # floats: every segment of the flat buffer starts 256-byte aligned


def _segments(sizes):
    offs, total = [], 0
    for n in sizes:
        offs.append(total)
        total += (n + _ALIGN - 1) // _ALIGN * _ALIGN
    return offs, total


def _flat_buffer(total):
    from .cuda_array import CudaArray
    flat = _state['flat']
    if flat is None or flat.size < total:
        flat = CudaArray.empty((total,))
        from .ops import L
        L().mc_memset(flat.ptr, 0, flat.nbytes)
        _state['flat'] = flat
        _state['flat_t'] = torch_view(flat)
    return flat, _state['flat_t']


def _raw_grads(grads):
    from minpy import array as marray
    raws = []
    for g in grads:
        if isinstance(g, marray.Array):
            raws.append(g._get_latest_data())  # pylint: disable=protected-access
        elif isinstance(g, marray.Number):
            raws.append(float(g.val))
        else:
            raws.append(g)
    return raws


def pack_gradients(grads):
    """Device gradients -> ONE flat buffer (the multi-tensor pack kernel) + zero-copy views of it in
    the form of `grads`.  Returns (flat, views); flat is None when no gradient lives on the device."""
    from minpy import array as marray
    from .cuda_array import CudaArray
    raws = _raw_grads(grads)
    if not any(isinstance(r, CudaArray) for r in raws):
        return None, grads
    from . import ops
    lib = ops.L()
    shapes = [r.shape if isinstance(r, (CudaArray, np.ndarray)) else None for r in raws]
    offs, total = _segments([ops._prod(s) if s is not None else 0 for s in shapes])  # pylint: disable=protected-access
    # a fresh flat buffer per call: the views handed back stay valid for as long as the
    # caller keeps them (the pool recycles the block as soon as they die)
    window = _state.get('p2p')
    if window is not None and 4 * total <= window[1] and _state['world'] > 1:
        # the bucket lives in the peer-memory window (same offset on every rank): mc_allreduce_avg then
        # runs the two-shot peer all-reduce.  The views stay valid until the next step packs again.
        flat = CudaArray(_WindowBuffer(window[0], window[1]), window[0], (total,))
    else:
        flat = CudaArray.empty((total,))
    srcs, dsts, ns = [], [], []
    for r, off, shape in zip(raws, offs, shapes):
        if shape is None:
            continue
        if isinstance(r, np.ndarray):
            r = CudaArray.from_numpy(r)
        r = ops.as_f32(r)
        srcs.append(r.ptr)
        dsts.append(flat.ptr + 4 * off)
        ns.append(r.size)
    k = len(srcs)
    if k:
        lib.mc_multi_copy((ctypes.c_void_p * k)(*srcs), (ctypes.c_void_p * k)(*dsts), (ctypes.c_int64 * k)(*ns), k)
    out = []
    for r, off, shape in zip(raws, offs, shapes):
        if shape is None:
            out.append(r)
        else:
            out.append(marray.wrap(CudaArray(flat.buf, flat.ptr + 4 * off, tuple(shape))))
    return flat, out


def reduce_flat(flat, view=None):
    """In-place average of a flat device buffer over all ranks (NCCL on torch's stream, which is
    ordered with the library's blocking stream through the legacy default stream)."""
    if _state['world'] <= 1:
        return
    if os.environ.get('MINPY_DP_DEBUG_SKIP_ALLREDUCE') == '1':   # timing experiments only: wrong gradients
        return
    if _state['native']:
        from .ops import L
        L().mc_allreduce_avg(flat.ptr, flat.size)
        return
    if _state['backend'] != 'nccl':   # pragma: no cover  (device arrays with a host backend)
        raise RuntimeError('device gradients need the nccl backend')
    dist = _dist()
    dist.all_reduce(view if view is not None else torch_view(flat), op=dist.ReduceOp.AVG)


def allreduce_mean(grads):
    """Average a list of gradients (MinPy Arrays, raw arrays or the Python scalar 0.0 for
    parameters the loss does not reach) over all ranks; returns a list of the same form."""
    world = _state['world']
    if world <= 1:
        return grads
    from minpy import array as marray
    dist = _dist()
    flat, out = pack_gradients(grads)
    if flat is not None:
        reduce_flat(flat)
        return out
    raws = _raw_grads(grads)
    # host gradients (NumPy policy): used by the CPU tests of this logic and by the CPU baseline
    import torch
    host = [np.asarray(r, dtype=np.float64) for r in raws if isinstance(r, np.ndarray)]
    if not host:
        return grads
    flat = np.concatenate([h.reshape(-1) for h in host])
    t = torch.from_numpy(flat)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    flat /= world
    out, pos = [], 0
    for r in raws:
        if isinstance(r, np.ndarray):
            n = r.size
            out.append(marray.wrap(flat[pos:pos + n].reshape(r.shape).astype(r.dtype, copy=False)))
            pos += n
        else:
            out.append(r)
    return out


def broadcast_params(params, src=0):
    """Every rank takes rank `src`'s parameter values (dict name -> MinPy Array, in place in
    the dict).  Called once after `Solver.init`."""
    if _state['world'] <= 1:
        return
    import torch
    from minpy import array as marray
    from .cuda_array import CudaArray
    dist = _dist()
    for name in sorted(params.keys()):
        value = params[name]
        raw = value._get_latest_data() if isinstance(value, marray.Array) else value  # pylint: disable=protected-access
        if isinstance(raw, CudaArray):
            dist.broadcast(torch_view(raw), src=src)
        else:
            host = np.ascontiguousarray(raw)
            t = torch.from_numpy(host)
            dist.broadcast(t, src=src)
            params[name] = marray.wrap(host)


def broadcast_host_array(arr, src=0):
    """Rank `src`'s copy of a small host array (same shape and dtype on every rank), e.g. the shuffle
    order of NDArrayIter.  A no-op without a process group."""
    if _state['world'] <= 1:
        return arr
    box = [np.asarray(arr) if _state['rank'] == src else None]
    _dist().broadcast_object_list(box, src=src)
    return box[0]


def allreduce_scalar_sum(value):
    """Sum of a Python number over ranks (accuracy counts, loss for printing)."""
    if _state['world'] <= 1:
        return value
    import torch
    dist = _dist()
    dev = 'cuda' if _state['backend'] == 'nccl' else 'cpu'
    t = torch.tensor([float(value)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def allgather_int(value):
    """The integer every rank holds, as a list indexed by rank (checksums in bench.py's dp_parity)."""
    if _state['world'] <= 1:
        return [int(value)]
    out = [None] * _state['world']
    _dist().all_gather_object(out, int(value))
    return out


def max_over_ranks(value):
    """Max of a Python number over ranks (bench timing)."""
    if _state['world'] <= 1:
        return value
    import torch
    dist = _dist()
    dev = 'cuda' if _state['backend'] == 'nccl' else 'cpu'
    t = torch.tensor([float(value)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
