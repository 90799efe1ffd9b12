"""CUDA-graph replay of `Solver._step` (opt-in: `Solver(..., step_graph=True)` or MINPY_STEP_GRAPH=1).

A MinPy training step is define-by-run Python: ~20-2000 library calls per step, each costing
5-50 us of interpreter + ctypes time depending on the host (profiles/r1g vs r1j: the same CNN step
took 0.89 ms on one box and 2.03 ms on another with identical kernel time).  When the step is
static -- same shapes, same control flow, no host read of a device value inside it -- the whole
launch sequence is captured ONCE from the unchanged Python step and replayed with one
cudaGraphLaunch per step.  The reference has no counterpart (its MXNet engine is asynchronous
but still pays one engine push per primitive, minpy/primitive.py:236-240).

What makes a replay a NEW step instead of a repetition of the captured one:
  * inputs: the batch is copied into static device buffers the captured kernels read;
  * state: parameters, auxiliary parameters (BatchNorm running statistics) and optimizer state are
    rebound functionally by the Python step (solver.py:190-195); the graph ends with device copies
    of every new value back into the buffer the captured kernels read it from, and the Python
    containers keep pointing at those buffers;
  * randomness and adam's step count: kernels captured while `mc_graph_begin` is active read a
    device-side replay counter (csrc/runtime.cu `graph_epoch`), so dropout masks differ from replay
    to replay (distribution-equal to the eager stream, not draw-equal) and adam's bias correction
    advances;
  * hyper-parameters baked into kernel arguments (learning rate ...): a signature of every scalar in
    the optimizer configs is checked before each replay; a change (lr decay) triggers re-capture.
Data parallel (world > 1): the all-reduce of the flat gradient buffer is a node of the same graph when
the library's own NCCL communicator is up (csrc/comm.cu); with torch's communicator it runs between
two graphs.

Anything the capture cannot express (a device->host read inside the loss function, a host array
created inside the step, shapes that change) aborts the capture, restores the Python state and
the solver continues eagerly -- never silently wrong.
"""
from __future__ import annotations

import ctypes
import warnings

import numpy as np

from ._lib import get_lib
from .cuda_array import CudaArray, BufferPool, _Buffer


class StepGraphUnsupported(RuntimeError):
    """The step cannot be captured; the solver falls back to eager execution."""


def _raw(value):
    """CudaArray behind a MinPy Array / raw value, or None when it is not a device array."""
    from minpy import array as marray
    if isinstance(value, marray.Array):
        if not value.on_device():
            return None
        value = value._get_latest_data()  # pylint: disable=protected-access
    return value if isinstance(value, CudaArray) else None


def _dtype_on_device(arr):
    k = arr.dtype.kind
    if k in 'fb':
        return np.dtype('float32')
    if arr.dtype == np.dtype('int64'):
        return np.dtype('int64')
    if k in 'iu':
        return np.dtype('int32')
    raise StepGraphUnsupported('unsupported input dtype %s' % arr.dtype)


class _StaticInput(object):
    """A device buffer the captured kernels read one batch tensor from."""

    def __init__(self, source):
        from minpy import array as marray
        raw = _raw(source)
        if raw is not None:
            self.dev = CudaArray.empty(raw.shape, raw.dtype)
        else:
            host = source.asnumpy() if isinstance(source, marray.Array) else np.asarray(source)
            self.dev = CudaArray.empty(host.shape, _dtype_on_device(host))
        self.wrapped = marray.wrap(self.dev)
        self.fill(source)

    def matches(self, source):
        shape = tuple(source.shape)
        return shape == tuple(self.dev.shape)

    def fill(self, source):
        from minpy import array as marray
        lib = get_lib()
        raw = _raw(source)
        if raw is not None:
            if raw.dtype != self.dev.dtype:
                raise StepGraphUnsupported('input dtype changed')
            if raw.ptr != self.dev.ptr and self.dev.nbytes:
                lib.mc_memcpy_d2d(self.dev.ptr, raw.ptr, self.dev.nbytes)
            return 0
        host = source.get_data(0) if isinstance(source, marray.Array) else source   # ArrayType.NUMPY == 0
        host = np.ascontiguousarray(host, dtype=self.dev.dtype)
        if host.nbytes:
            lib.mc_memcpy_h2d(self.dev.ptr, host.ctypes.data, host.nbytes)
        self._keep = host          # a pinned source is copied asynchronously: keep it alive
        return host.nbytes


class StepGraph(object):
    """Capture-and-replay driver owned by one Solver."""

    def __init__(self, solver, warmup=2):
        self.solver = solver
        self.warmup = warmup          # eager steps first: optimizer state exists, caches are warm
        self.seen = 0
        self.disabled = False
        self.execs = None             # [graph before the all-reduce, graph after] or [whole step]
        self.sig = None
        self.inputs = None            # [[_StaticInput per data tensor], [per label tensor]]
        self.loss_raw = None
        self.flat = None              # flat gradient buffer (world > 1)
        self.flat_view = None
        self.private_pool = None      # device blocks reserved for the graph's temporaries
        self.adam_cfgs = []
        self.state_objs = []
        self.replays = 0
        self.captures = 0
        self.h2d_bytes = 0

    # ---- pieces ---------------------------------------------------------------------------
    def _state_entries(self):
        """(container, key, Array object, CudaArray) of everything a step rebinds."""
        solver = self.solver
        out = []
        for container in [solver.model.params, getattr(solver.model, 'aux_params', {})] + \
                [solver.optim_configs[k] for k in solver.model.params.keys()]:
            for key, value in container.items():
                raw = _raw(value)
                if raw is not None:
                    out.append((container, key, value, raw))
        return out

    def _hyper_signature(self):
        sig = []
        for name in self.solver.model.params.keys():
            cfg = self.solver.optim_configs[name]
            for key in sorted(cfg.keys()):
                v = cfg[key]
                if key != 't' and isinstance(v, (int, float, np.integer, np.floating)):
                    sig.append((name, key, float(v)))
        return tuple(sig)

    def _bind_inputs(self, batch):
        from minpy.nn.io import DataBatch
        groups = [list(batch.data), list(batch.label)]
        if self.inputs is None:
            self.inputs = [[_StaticInput(src) for src in g] for g in groups]
            self.execs = None
        else:
            if [len(g) for g in groups] != [len(g) for g in self.inputs]:
                return None
            for g, statics in zip(groups, self.inputs):
                for src, st in zip(g, statics):
                    if not st.matches(src):
                        return None
            for g, statics in zip(groups, self.inputs):
                for src, st in zip(g, statics):
                    self.h2d_bytes += st.fill(src)
        return DataBatch([st.wrapped for st in self.inputs[0]], [st.wrapped for st in self.inputs[1]],
                         pad=batch.pad, index=batch.index)

    def _drop(self):
        lib = get_lib()
        if self.execs:
            lib.mc_sync()
            for ex in self.execs:
                lib.mc_graph_destroy(ex)
        self.execs = None
        if self.private_pool is not None:
            self.private_pool.release()
            self.private_pool = None

    # ---- capture --------------------------------------------------------------------------
    def _capture(self, sbatch):
        from . import parallel
        solver = self.solver
        lib = get_lib()
        self._drop()
        lib.mc_sync()
        entries = self._state_entries()
        known = set((id(c), k) for c, k, _, _ in entries)
        scalars = [(cfg, dict((k, v) for k, v in cfg.items() if _raw(v) is None))
                   for cfg in solver.optim_configs.values()]
        # private pool: every temporary the captured kernels touch comes from blocks no eager
        # allocation can be handed later (a replay would overwrite live eager data otherwise)
        pool = BufferPool()
        _Buffer._active_pool = pool                            # pylint: disable=protected-access
        execs = []
        try:
            lib.mc_graph_begin()
            names, grads, loss = solver._step_grads(sbatch)   # pylint: disable=protected-access
            import os
            if parallel.world_size() > 1 or os.environ.get('MINPY_DP_DEBUG_FORCE_PACK') == '1':
                flat, grads = parallel.pack_gradients(grads)
                if flat is None:
                    raise StepGraphUnsupported('gradients are not on the device')
                self.flat = flat
                if parallel.native_comm() or parallel.world_size() <= 1:
                    parallel.reduce_flat(flat)            # NCCL on the library stream: captured
                else:
                    ex = ctypes.c_void_p()
                    lib.mc_graph_end(ctypes.byref(ex))
                    execs.append(ex)
                    lib.mc_graph_begin()
            solver._step_update(names, grads)                 # pylint: disable=protected-access
            if not getattr(solver, '_update_was_multi_tensor', False) and \
                    any('t' in cfg for cfg in solver.optim_configs.values()):
                # the per-parameter mc_adam bakes its bias correction from the host-side t at launch:
                # a replay would reuse the capture-time value.  Only mc_multi_update reads the replay counter.
                raise StepGraphUnsupported('a step-counting update rule ran outside the multi-tensor kernel')
            loss_raw = _raw(loss)
            if loss_raw is None:
                raise StepGraphUnsupported('the loss is not a device scalar')
            for container, key, old_obj, old_raw in entries:
                new_raw = _raw(container[key])
                if new_raw is None or new_raw.shape != old_raw.shape or new_raw.dtype != old_raw.dtype:
                    raise StepGraphUnsupported('state %r changed shape, dtype or placement' % (key,))
                if new_raw.ptr != old_raw.ptr and old_raw.nbytes:
                    lib.mc_memcpy_d2d(old_raw.ptr, new_raw.ptr, old_raw.nbytes)
                container[key] = old_obj
            for container, key, _, _ in self._state_entries():
                if (id(container), key) not in known:
                    raise StepGraphUnsupported('state %r was created inside the captured step' % (key,))
            ex = ctypes.c_void_p()
            lib.mc_graph_end(ctypes.byref(ex))
            execs.append(ex)
        except Exception as exc:  # pylint: disable=broad-except
            lib.cdll.mc_graph_abort()
            for ex in execs:
                lib.cdll.mc_graph_destroy(ex)
            for container, key, old_obj, _ in entries:
                container[key] = old_obj
            for cfg, saved in scalars:
                cfg.update(saved)
            self.disabled = True
            _Buffer._active_pool = None                        # pylint: disable=protected-access
            pool.release()
            warnings.warn('step graph disabled, continuing eagerly: %s' % (exc,))
            return False
        finally:
            _Buffer._active_pool = None                        # pylint: disable=protected-access
        self.private_pool = pool
        self.execs = execs
        self.state_objs = [obj for _, _, obj, _ in entries]
        self.loss_raw = loss_raw
        self.flat_view = parallel.torch_view(self.flat) if len(execs) == 2 else None
        self.adam_cfgs = [cfg for cfg in solver.optim_configs.values() if 't' in cfg]
        self.sig = self._hyper_signature()
        self.captures += 1
        # the capture pass advanced adam's host-side t once without executing anything: the first
        # replay is that step
        self._launch(first=True)
        return True

    def _launch(self, first=False):
        from . import parallel
        from minpy import array as marray
        lib = get_lib()
        lib.mc_graph_launch(self.execs[0])
        if len(self.execs) == 2:
            parallel.reduce_flat(self.flat, self.flat_view)
            lib.mc_graph_launch(self.execs[1])
        if not first:
            for cfg in self.adam_cfgs:
                cfg['t'] = int(cfg['t']) + 1
        self.replays += 1
        if self.replays % 64 == 0 and parallel.p2p_active():
            # the peer all-reduce kernel gives up on a missing rank after its watchdog (csrc/comm.cu) and
            # flags it in host-visible memory; replays launch the captured kernel directly, so look here
            ptr, size, bad = ctypes.c_void_p(), ctypes.c_size_t(), ctypes.c_int()
            lib.mc_p2p_info(ctypes.byref(ptr), ctypes.byref(size), ctypes.byref(bad))
            if bad.value:
                raise RuntimeError('peer-memory all-reduce timed out waiting for a rank: gradients of the last '
                                   'steps were not reduced')
        # every replay rewrites parameters / running statistics / optimizer state IN PLACE: a host copy an
        # earlier asnumpy() cached inside the Array is stale from here on (minpy/array.py get_data)
        for obj in self.state_objs:
            if isinstance(obj, marray.Array) and len(obj._data) > 1:   # pylint: disable=protected-access
                obj._data = {2: obj._data[2]}                           # pylint: disable=protected-access
                obj._latest_version = 2                                 # ArrayType.CUDA
        keep = CudaArray.empty(())
        lib.mc_memcpy_d2d(keep.ptr, self.loss_raw.ptr, 4)
        keep.begin_readback()          # the loss travels to the host behind this step, on a third stream
        self.solver.loss_history.append(marray.wrap(keep))

    def writes_in_place(self):
        """True while replays update the solver's state inside its existing buffers (callers that keep
        references to old parameter values -- Solver.best_params -- must copy instead)."""
        return self.execs is not None and not self.disabled

    # ---- the step -------------------------------------------------------------------------
    def step(self, batch):
        solver = self.solver
        if self.disabled:
            return solver._step_eager(batch)                  # pylint: disable=protected-access
        if self.seen < self.warmup:
            self.seen += 1
            return solver._step_eager(batch)                  # pylint: disable=protected-access
        try:
            sbatch = self._bind_inputs(batch)
        except StepGraphUnsupported as exc:
            self.disabled = True
            warnings.warn('step graph disabled, continuing eagerly: %s' % (exc,))
            return solver._step_eager(batch)                  # pylint: disable=protected-access
        if sbatch is None:                                    # a batch of another shape: run it eagerly
            return solver._step_eager(batch)                  # pylint: disable=protected-access
        if self.execs is None or self._hyper_signature() != self.sig:
            if not self._capture(sbatch):
                return solver._step_eager(batch)              # pylint: disable=protected-access
            return None
        self._launch()
        return None
