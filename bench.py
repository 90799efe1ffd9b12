#!/usr/bin/env python
"""bench.py -- training samples/sec of the MinPy hot path on the B200 `cuda` array variant.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload mlp_c3|mlp_c1|cnn_c4|cnn_ref|lstm_c5]
                    [--impl b200|reference] [--no-ufunc] [--no-cpu-baseline]

One "step" is one `Solver._step` (forward + tape backward + gradient all-reduce when N > 1 +
one update rule call per parameter) over one minibatch of synthetic CIFAR-shaped data.  Rank 0
prints ONE JSON line (contract: see DESIGN.md "Measurement").

* `value`  : samples/s with the batches already resident in HBM, CUDA events on the library's
             stream, barrier + synchronize on both sides, max over ranks.
* `e2e`    : the same loop fed from PINNED HOST memory through `NDArrayIter` (one H2D copy of
             the batch per step) and reading the loss back every step (D2H).
* `roofline`: the dominant kernel family of the step, timed with CUDA events around every
             launch in a second, instrumented pass of the same steps.
* `ufunc`  : GB/s of the elementwise / broadcast / reduction kernels (BASELINE config 2).
* `cpu_baseline`: the NumPy oracle port of the reference's path, timed on this box's cores.

`--impl reference` times the reference's CPU path (the oracle port, float64 as the
reference's NumPy variant computes) on the same workload and prints the same line shape.
"""
from __future__ import annotations

import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = 'training samples/sec'
L2_BYTES = 126e6


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# ------------------------------------------------------------------------------ clocks ----
class ClockSampler(object):
    """Samples SM clocks and throttle reasons with nvidia-smi while the timed region runs."""

    FIELDS = ('clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
              'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, device):
        self.device = device
        self.samples = []
        self.proc = None
        self.thread = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ['nvidia-smi', '-i', str(self.device), '--query-gpu=' + self.FIELDS, '--format=csv,noheader,nounits',
                 '-lms', '100'], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return
        self.thread = threading.Thread(target=self._read, daemon=True)
        self.thread.start()

    def _read(self):
        for line in self.proc.stdout:
            parts = [p.strip() for p in line.split(',')]
            if len(parts) >= 6:
                self.samples.append(parts)

    def stop(self):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:  # pylint: disable=broad-except
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for parts in self.samples:
            try:
                sm.append(float(parts[0]))
                mx = float(parts[1])
            except ValueError:
                continue
            for name, flag in zip(names, parts[2:6]):
                if flag.lower().startswith('active'):
                    reasons.add(name)
        return {'sm_mhz': float(np.median(sm)) if sm else None, 'sm_max_mhz': mx, 'reasons': sorted(reasons),
                'samples': len(sm)}


# --------------------------------------------------------------------- per-call timing ----
class CallTimer(object):
    """CUDA-event pair around every library call (installed on the ctypes binding) so that the
    kernel families of one step can be ranked and the dominant one put on the roofline."""

    def __init__(self, lib):
        self.lib = lib
        self.records = []          # (name, start_event, stop_event, work) ; work = (flops, bytes)
        self.pool = []
        self.calls = []            # (name, raw function, args) in launch order, for replay_family

    def _event(self):
        if self.pool:
            return self.pool.pop()
        ev = ctypes.c_void_p()
        self.lib.mc_event_create(ctypes.byref(ev))
        return ev

    def install(self):
        lib = self.lib
        self.saved = {}
        skip = ('mc_event', 'mc_alloc', 'mc_free', 'mc_sync', 'mc_launch_count', 'mc_last_error', 'mc_get', 'mc_set',
                'mc_init', 'mc_host', 'mc_mem_info', 'mc_device_count', 'mc_sm_count', 'mc_graph', 'mc_gemm_last_path', 'mc_side_')
        for name in lib.decls:
            if name.startswith(skip):
                continue
            fn = getattr(lib, name)
            self.saved[name] = fn
            setattr(lib, name, self._wrap(name, fn))

    def _wrap(self, name, fn):
        def call(*args):
            e0, e1 = self._event(), self._event()
            self.lib.mc_event_record(e0)
            fn(*args)
            self.lib.mc_event_record(e1)
            self.records.append((name, e0, e1, _work_of(name, args)))
            self.calls.append((name, fn, args))
        return call

    def replay_family(self, family, steps_recorded, reps=20):
        """Average device time of ONE step's worth of `family` calls, launched back to back from a
        CUDA graph (no host gaps between kernels), CUDA events on the library stream.  The calls are
        the recorded ones with their recorded arguments; their outputs land in pool memory nobody
        reads afterwards (this runs last)."""
        lib = self.lib
        calls = [(fn, args) for name, fn, args in self.calls if name == family]
        per_step = len(calls) // max(1, steps_recorded)
        calls = calls[-per_step:] if per_step else []
        if not calls:
            return None
        try:
            lib.mc_sync()
            lib.mc_graph_begin()
            for fn, args in calls:
                fn(*args)
            ex = ctypes.c_void_p()
            lib.mc_graph_end(ctypes.byref(ex))
        except Exception as exc:  # pylint: disable=broad-except
            lib.cdll.mc_graph_abort()
            log('replay of %s not possible: %s' % (family, exc))
            return None
        for _ in range(3):
            lib.mc_graph_launch(ex)
        e0, e1 = self._event(), self._event()
        lib.mc_sync()
        lib.mc_event_record(e0)
        for _ in range(reps):
            lib.mc_graph_launch(ex)
        lib.mc_event_record(e1)
        lib.mc_sync()
        ms = ctypes.c_float()
        lib.mc_event_elapsed_ms(e0, e1, ctypes.byref(ms))
        lib.mc_graph_destroy(ex)
        return {'calls': len(calls), 'ms_per_step': ms.value / reps}

    def uninstall(self):
        for name, fn in self.saved.items():
            setattr(self.lib, name, fn)

    def summary(self):
        self.lib.mc_sync()
        fam = {}
        ms = ctypes.c_float()
        for name, e0, e1, work in self.records:
            self.lib.mc_event_elapsed_ms(e0, e1, ctypes.byref(ms))
            key = name if name != 'mc_gemm' else 'mc_gemm'
            f = fam.setdefault(key, {'calls': 0, 'ms': 0.0, 'flops': 0.0, 'bytes': 0.0})
            f['calls'] += 1
            f['ms'] += ms.value
            f['flops'] += work[0]
            f['bytes'] += work[1]
            self.pool.extend((e0, e1))
        self.records = []
        return fam


def _i(v):
    return int(v.value) if hasattr(v, 'value') else int(v)


def _work_of(name, a):
    """(algorithmic flops, algorithmic bytes) of one library call: compulsory traffic, every
    input read once and every output written once."""
    try:
        if name == 'mc_gemm':
            M, N, K = _i(a[2]), _i(a[3]), _i(a[4])
            return 2.0 * M * N * K, 4.0 * (M * K + K * N + M * N)
        if name == 'mc_unary':
            return 0.0, 8.0 * _i(a[3])
        if name == 'mc_binary':
            return 0.0, 12.0 * _i(a[4])
        if name == 'mc_binary_scalar':
            return 0.0, 8.0 * _i(a[5])
        if name == 'mc_binary_vec':
            return 0.0, 8.0 * _i(a[6]) * _i(a[7])
        if name == 'mc_reduce':
            return 0.0, 4.0 * (_i(a[3]) * _i(a[4]) * _i(a[5]) + _i(a[3]) * _i(a[5]))
        if name == 'mc_batchnorm_train_fwd':
            N, D = _i(a[10]), _i(a[11])
            return 0.0, 4.0 * (2 * N * D + 6 * D)
        if name == 'mc_batchnorm_train_bwd':
            N, D = _i(a[8]), _i(a[9])
            return 0.0, 4.0 * (3 * N * D + 4 * D)
        if name == 'mc_bn_dropout_fwd':
            N, D = _i(a[10]), _i(a[11])
            return 0.0, 4.0 * (2 * N * D + 6 * D)            # batchnorm fwd with the dropout mask folded in
        if name == 'mc_bn_dropout_relu_bwd':
            N, D = _i(a[9]), _i(a[10])
            return 0.0, 4.0 * (3 * N * D + 5 * D)            # x, dy in; dx out; dgamma, dbeta, dbias
        if name == 'mc_head_softmax_loss':
            r, K, C = _i(a[-3]), _i(a[-2]), _i(a[-1])
            return 4.0 * r * K * C, 4.0 * (2 * r * K + K * C + 3 * r * C)
        if name == 'mc_softmax_ce_fused':
            r, c = _i(a[-2]), _i(a[-1])
            return 0.0, 8.0 * r * c
        if name == 'mc_dropout':
            return 0.0, 8.0 * _i(a[2])
        if name == 'mc_vjp_chain':
            return 0.0, 8.0 * _i(a[8])
        if name in ('mc_softmax_ce_fwd', 'mc_softmax_ce_bwd', 'mc_softmax_rows'):
            r, c = _i(a[-2]), _i(a[-1])
            return 0.0, 8.0 * r * c
        if name == 'mc_sgd_momentum':
            return 0.0, 20.0 * _i(a[7])
        if name == 'mc_adam':
            return 0.0, 28.0 * _i(a[12])
        if name in ('mc_memcpy_d2d', 'mc_memset'):
            return 0.0, (2.0 if name == 'mc_memcpy_d2d' else 1.0) * _i(a[2])
        if name == 'mc_conv2d_nhwc_fwd_pool':
            N, C, H, W, Cp, wpad, F, Fp, KH, KW, pad = [_i(v) for v in a[5:16]]
            OH, OW = H + 2 * pad - KH + 1, W + 2 * pad - KW + 1
            return (2.0 * N * OH * OW * F * C * KH * KW,
                    4.0 * (N * H * W * Cp + F * C * KH * KW) + 5.0 * N * (OH // 2) * (OW // 2) * Fp)
        if name in ('mc_conv2d_nhwc_fwd', 'mc_conv2d_nhwc_wgrad'):
            N, C, H, W, Cp, wpad, F, Fp, KH, KW, pad = [_i(v) for v in a[4:15]]
            OH, OW = H + 2 * pad - KH + 1, W + 2 * pad - KW + 1
            return (2.0 * N * OH * OW * F * C * KH * KW,
                    4.0 * (N * H * W * Cp + N * OH * OW * Fp + F * C * KH * KW))
        if name == 'mc_conv2d_nhwc_dgrad':
            N, C, H, W, Cp, F, Fp, KH, KW, pad = [_i(v) for v in a[3:13]]
            OH, OW = H + 2 * pad - KH + 1, W + 2 * pad - KW + 1
            return (2.0 * N * OH * OW * F * C * KH * KW,
                    4.0 * (N * H * W * Cp + N * OH * OW * Fp + F * C * KH * KW))
        if name == 'mc_maxpool2d_nhwc_fwd':
            N, H, W, Cp, KH, KW, st = [_i(v) for v in a[3:10]]
            OH, OW = (H - KH) // st + 1, (W - KW) // st + 1
            return 0.0, 4.0 * N * H * W * Cp + 5.0 * N * OH * OW * Cp
        if name == 'mc_maxpool2d_nhwc_bwd':
            N, H, W, Cp, K, OH, OW = [_i(v) for v in a[4:11]]
            return 0.0, 4.0 * N * H * W * Cp + 9.0 * N * OH * OW * Cp
        if name in ('mc_nchw_to_nhwc', 'mc_nhwc_to_nchw'):
            N, C, H, W = [_i(v) for v in a[2:6]]
            return 0.0, 8.0 * N * C * H * W
        if name == 'mc_multi_update':
            k = _i(a[1])
            per = {0: 12.0, 1: 20.0, 2: 20.0, 3: 28.0}[_i(a[0])]
            return 0.0, per * sum(int(a[9][j]) for j in range(k))
        if name == 'mc_lstm_cell_fwd':
            N, H = _i(a[5]), _i(a[6])
            return 0.0, 4.0 * N * H * (4 + 1 + 2 + 4)
        if name == 'mc_lstm_cell_bwd':
            N, H = _i(a[7]), _i(a[8])
            return 0.0, 4.0 * N * H * (4 + 2 + 2 + 4 + 1)
        if name in ('mc_lstm_seq_fwd', 'mc_lstm_seq_bwd'):
            # per time step: the recurrent GEMM (N,H)x(H,4H) [bwd: (N,4H)x(4H,H)] with Wh read once, plus the
            # gate kernel's operands (fwd: a, c in; h, c, gates out; bwd: gates, c, c', dh, dc in; da, dc, dh out)
            T, N, H = _i(a[5]), _i(a[6]), _i(a[7])
            per_nh = 12 if name == 'mc_lstm_seq_fwd' else 14
            return T * 2.0 * N * H * 4 * H, T * 4.0 * (per_nh * N * H + 4 * H * H)
    except Exception:  # pylint: disable=broad-except
        pass
    return 0.0, 0.0


# ---------------------------------------------------------------------------- peaks -------
def measured_peaks():
    path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return {'hbm_gbs': p['hbm_gbs'], 'bf16_tflops': p['bf16_tflops'],
                'bf16_tflops_sustained': p.get('bf16_tflops_sustained', p['bf16_tflops']), 'source': 'measured'}
    return {'hbm_gbs': 6650.0, 'bf16_tflops': 1590.0, 'bf16_tflops_sustained': 1400.0, 'source': 'fallback'}


def ncu_traffic(workload, family):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `family` from the committed
    `ncu --set full` capture (profiles/traffic.json, written by profiles/summarize_ncu.py), or None."""
    path = os.path.join(ROOT, 'profiles', 'traffic.json')
    try:
        with open(path) as f:
            return json.load(f).get(workload, {}).get(family)
    except (OSError, ValueError):
        return None


# ---------------------------------------------------------------- the b200 training arm ----
def pinned_array(lib, shape, dtype):
    """NumPy array over pinned host memory from the library (kept alive by the returned base)."""
    dtype = np.dtype(dtype)
    n = int(np.prod(shape))
    p = ctypes.c_void_p()
    lib.mc_host_alloc(ctypes.byref(p), n * dtype.itemsize)
    buf = (ctypes.c_char * (n * dtype.itemsize)).from_address(p.value)
    arr = np.frombuffer(buf, dtype=dtype).reshape(shape)
    return arr


def bench_config(workload, world, nbatch):
    """The `config` object of the JSON line: the same for the b200 and the reference arm."""
    from minpy_b200 import workloads
    _, batch, _, _, make_data, desc = workloads.WORKLOADS[workload]
    x, y = make_data(1, seed=0)
    in_bytes = batch * (x.nbytes + y.nbytes)
    return {'workload': workload, 'description': desc, 'per_gpu_batch': batch, 'global_batch': batch * world,
            'parallelism': 'dp%d' % world,
            'l2': '%d distinct batches/rank (%.0f MB) cycled, larger than the 126 MB L2' % (nbatch, nbatch * in_bytes / 1e6)}


def build_solver(workload, train_iter, step_graph=False):
    from minpy.nn.solver import Solver
    from minpy_b200 import workloads
    factory, _, rule, optim_config, _, _ = workloads.WORKLOADS[workload]
    model = factory()
    solver = Solver(model, train_iter, train_iter, update_rule=rule, optim_config=dict(optim_config),
                    init_rule='gaussian', init_config={'stdvar': 0.01}, verbose=False, step_graph=step_graph)
    solver.init()
    return model, solver


def run_b200(args):
    import minpy
    minpy.set_global_policy('only_cuda')
    import minpy.numpy as mnp
    from minpy import array as marray
    from minpy.nn.io import NDArrayIter, DataBatch
    from minpy_b200 import parallel, workloads
    from minpy_b200._lib import get_lib
    from minpy_b200.cuda_array import ensure_device, CudaArray

    world = parallel.init()
    rank = parallel.rank()
    if world != args.gpus:
        log('note: --gpus %d but WORLD_SIZE=%d; using WORLD_SIZE' % (args.gpus, world))
    device = ensure_device()
    lib = get_lib()
    _, batch, _, _, make_data, desc = workloads.WORKLOADS[args.workload]

    # distinct batches per rank: more bytes than the 126 MB L2 so that no step re-reads its
    # inputs from cache (weights stay L2-resident, as they do in real training)
    nbatch = args.pool
    X, Y = make_data(batch * nbatch, seed=100 + rank)
    in_bytes = X[:batch].nbytes + Y[:batch].nbytes
    Xp = pinned_array(lib, X.shape, X.dtype)
    Yp = pinned_array(lib, Y.shape, Y.dtype)
    Xp[...] = X
    Yp[...] = Y
    host_iter = NDArrayIter(Xp, Yp, batch_size=batch, shuffle=False, num_parts=1, part_index=0,
                            device_prefetch=bool(args.prefetch))
    model, solver = build_solver(args.workload, host_iter, step_graph=bool(args.graph))

    # device-resident batches for the kernel-only leg
    dev_batches = []
    for i in range(nbatch):
        xb = marray.wrap(CudaArray.from_numpy(X[i * batch:(i + 1) * batch]))
        yb = marray.wrap(CudaArray.from_numpy(Y[i * batch:(i + 1) * batch]))
        dev_batches.append(DataBatch([xb], [yb], pad=0, index=i))
    lib.mc_sync()

    def host_batches():
        while True:
            host_iter.reset()
            for b in host_iter:
                yield b
    hb = host_batches()

    def step_device(i):
        solver._step(dev_batches[i % nbatch])  # pylint: disable=protected-access

    def step_e2e(_i):
        solver._step(next(hb))  # pylint: disable=protected-access
        # D2H read of a step's result, every step: each loss travels to pinned host memory behind its
        # step (mc_readback_begin, third stream); the host reads it ONE STEP BEHIND, i.e. it waits for
        # step i-1 while step i runs (the reference blocks on step i itself, solver.py:187)
        hist = solver.loss_history
        return float(hist[-2]) if len(hist) >= 2 else None

    def step_e2e_blocking(_i):
        solver._step(next(hb))  # pylint: disable=protected-access
        return float(solver.loss_history[-1])          # the reference's pattern: wait for this very step

    def finish_e2e():
        return float(solver.loss_history[-1])          # the last step's loss, inside the timed region

    def window(step_fn, steps, first, finish=None):
        """EXACTLY `steps` steps between two events, barrier + synchronize on both sides."""
        lib.mc_sync()
        parallel.barrier()
        lib.mc_sync()
        e0, e1 = ctypes.c_void_p(), ctypes.c_void_p()
        lib.mc_event_create(ctypes.byref(e0))
        lib.mc_event_create(ctypes.byref(e1))
        n0, n1 = ctypes.c_uint64(), ctypes.c_uint64()
        lib.mc_launch_count(ctypes.byref(n0))
        t0 = time.perf_counter()
        lib.mc_event_record(e0)
        for i in range(steps):
            step_fn(first + i)
        if finish is not None:
            finish()
        lib.mc_event_record(e1)
        lib.mc_sync()
        parallel.barrier()
        wall = time.perf_counter() - t0
        lib.mc_launch_count(ctypes.byref(n1))
        ms = ctypes.c_float()
        lib.mc_event_elapsed_ms(e0, e1, ctypes.byref(ms))
        lib.mc_event_destroy(e0)
        lib.mc_event_destroy(e1)
        return ms.value, n1.value - n0.value, wall * 1e3

    def timed(step_fn, steps, warmup, finish=None, min_ms=0.0):
        """W warm-up steps, then windows of EXACTLY `steps` steps each (CUDA events on the library
        stream, max over ranks per window) until `min_ms` of device time has been covered: one
        scheduler hiccup cannot move the figure.  Returns the MEDIAN window."""
        for i in range(warmup):
            step_fn(i)
        solver.loss_history = []
        wins, launches, walls, own = [], 0, [], []
        first = warmup
        while True:
            ms, launches, wall = window(step_fn, steps, first, finish)
            first += steps
            own.append(ms)
            wins.append(parallel.max_over_ranks(ms))
            walls.append(parallel.max_over_ranks(wall))
            if sum(wins) >= min_ms or len(wins) >= args.max_windows:
                break
            del solver.loss_history[:-2]
        if world > 1:
            log('rank %d: %s %.4f ms/step on its own device (median of %d windows)' % (
                rank, step_fn.__name__, float(np.median(own)) / steps, len(own)))
        mid = int(np.argsort(wins)[len(wins) // 2])
        stats = {'windows': len(wins), 'steps_per_window': steps, 'stat': 'median window',
                 'min_ms_per_step': round(min(wins) / steps, 4), 'max_ms_per_step': round(max(wins) / steps, 4),
                 'timed_ms_total': round(sum(wins), 2)}
        return wins[mid], launches, walls[mid], stats

    sampler = ClockSampler(device)
    if rank == 0:
        sampler.start()
    dev_ms, launches, dev_wall, dev_stats = timed(step_device, args.steps, args.warmup, min_ms=args.min_ms)
    e2e_ms, _, e2e_wall, e2e_stats = timed(step_e2e, args.steps, args.warmup, finish=finish_e2e, min_ms=args.min_ms)
    e2e_blocking_ms, _, _, _ = timed(step_e2e_blocking, args.steps, args.warmup)
    clocks = sampler.stop() if rank == 0 else None
    last_loss = float(solver.loss_history[-1])
    graph_info = None
    sg = getattr(solver, '_step_graph', None)
    if sg is not None:
        # the same steps launched eagerly from Python (one library call per kernel): the launch count
        # per step (= kernel nodes each replay executes) and the host-bound time it replaces
        graph_info = {'replays': sg.replays, 'captures': sg.captures, 'fell_back_to_eager': bool(sg.disabled)}
        was = sg.disabled
        sg.disabled = True
        eager_ms, launches, _, _ = timed(step_device, args.steps, args.warmup)
        sg.disabled = was
        graph_info['eager_ms_per_step'] = round(eager_ms / args.steps, 4)

    # instrumented pass: event pair around every library call of the same steps
    if sg is not None:
        sg.disabled = True
    timer = CallTimer(lib)
    timer.install()
    try:
        for i in range(min(args.steps, 20)):
            step_device(i)
        fam = timer.summary()
    finally:
        timer.uninstall()
    inst_steps = min(args.steps, 20)

    total_samples = batch * world * args.steps
    value = total_samples / (dev_ms * 1e-3)
    e2e_value = total_samples / (e2e_ms * 1e-3)
    peaks = measured_peaks()

    out = None
    if rank == 0:
        tot = sum(f['ms'] for f in fam.values()) or 1.0
        ranked = sorted(fam.items(), key=lambda kv: -kv[1]['ms'])
        log('kernel families over %d instrumented steps (event time, share of summed kernel time):' % inst_steps)
        for name, f in ranked[:14]:
            log('  %-26s calls/step %6.1f  us/call %8.2f  share %5.1f%%  %8.1f GB/s  %8.2f TFLOP/s' % (
                name, f['calls'] / inst_steps, 1e3 * f['ms'] / f['calls'], 100 * f['ms'] / tot,
                f['bytes'] / (f['ms'] * 1e-3) / 1e9 if f['ms'] else 0, f['flops'] / (f['ms'] * 1e-3) / 1e12 if f['ms'] else 0))
        # every family re-timed WITHOUT host gaps (its calls of one step replayed back to back from a graph)
        replayed = {}
        if not args.no_family_replay:
            for name, f in ranked[:10]:
                if name.startswith(('mc_allreduce', 'mc_broadcast', 'mc_readback', 'mc_memcpy_h2d', 'mc_prefetch')):
                    continue
                r = timer.replay_family(name, inst_steps)
                if r is not None:
                    replayed[name] = r
                    log('  replay %-26s %3d calls/step  %8.2f us/step  %7.2f us/call' % (
                        name, r['calls'], 1e3 * r['ms_per_step'], 1e3 * r['ms_per_step'] / max(1, r['calls'])))
            log('  replayed families sum to %.1f us/step; the timed step is %.1f us' % (
                1e3 * sum(r['ms_per_step'] for r in replayed.values()), 1e3 * dev_ms / args.steps))
        # the roofline is about the dominant COMPUTE kernel: the collective's event time is mostly the wait
        # for the slowest rank (it stays in kernel_families)
        compute = [kv for kv in ranked if not kv[0].startswith(('mc_allreduce', 'mc_broadcast'))] or ranked
        top_name, top = compute[0]
        # duration of the dominant family: its recorded calls of one step, replayed back to back from a
        # CUDA graph (the event pairs above include the host gap in front of every small kernel)
        rep = None if top_name.startswith(('mc_allreduce', 'mc_broadcast')) else (replayed.get(top_name) or timer.replay_family(top_name, inst_steps))
        calls_per_step = top['calls'] / inst_steps
        if rep is not None:
            fam_ms = rep['ms_per_step']
            timing = 'graph replay of the %d %s calls of one step, 20 repetitions, CUDA events' % (rep['calls'], top_name)
        else:
            fam_ms = top['ms'] / inst_steps
            timing = 'CUDA-event pair around every call'
        flops, nbytes = top['flops'] / inst_steps, top['bytes'] / inst_steps
        tf32_peak = peaks['bf16_tflops_sustained'] / 2.0          # kind::tf32 issues at half the bf16 rate
        t_tensor = flops / (tf32_peak * 1e12)
        t_hbm = nbytes / (peaks['hbm_gbs'] * 1e9)
        traffic = ncu_traffic(args.workload, top_name)
        if flops > 0 and t_tensor >= t_hbm:
            achieved = flops / (fam_ms * 1e-3) / 1e12
            peak = peaks['bf16_tflops_sustained']
            roofline = {'kernel': top_name + ' (tcgen05 kind::tf32)', 'bound': 'tensor', 'achieved': round(achieved, 2),
                        'peak': peak, 'unit': 'TFLOP/s', 'frac': round(achieved / peak, 4), 'traffic': traffic,
                        'frac_of_tf32_rate': round(achieved / tf32_peak, 4),
                        'peak_source': peaks['source'] + ' bf16 cuBLAS sustained (MEASURED_PEAKS.json); kind::tf32 runs '
                                       'at half that rate, frac_of_tf32_rate uses the halved figure'}
        else:
            achieved = nbytes / (fam_ms * 1e-3) / 1e9
            roofline = {'kernel': top_name, 'bound': 'hbm', 'achieved': round(achieved, 1), 'peak': peaks['hbm_gbs'],
                        'unit': 'GB/s', 'frac': round(achieved / peaks['hbm_gbs'], 4), 'traffic': traffic,
                        'peak_source': peaks['source'] + ' copy bandwidth (MEASURED_PEAKS.json)'}
            if flops > 0:
                roofline['tflops'] = round(flops / (fam_ms * 1e-3) / 1e12, 2)
        roofline.update({
            'avg_us_per_launch': round(1e3 * fam_ms / max(1.0, calls_per_step), 2),
            'launches_per_step': round(calls_per_step, 1),
            'algorithmic_flops_per_step': flops, 'algorithmic_bytes_per_step': nbytes,
            'arith_intensity_flop_per_byte': round(flops / max(1.0, nbytes), 1),
            'share_of_kernel_time': round(top['ms'] / tot, 3),
            'ideal_us_per_step': round(1e6 * max(t_tensor, t_hbm), 2), 'timing': timing})
        roofline['how'] = ('families ranked by a CUDA-event pair around every library call in an eager pass of %d '
                           'steps (summed %.3f ms/step, host gaps included) vs %.3f ms/step for the timed steps; '
                           'the dominant family is then re-timed without host gaps (see timing)' % (
                               inst_steps, tot / inst_steps, dev_ms / args.steps))
        out = {
            'metric': METRIC, 'value': round(value, 1), 'unit': 'samples/s', 'n_gpus': world, 'steps': args.steps,
            'warmup': args.warmup, 'ms_per_step': round(dev_ms / args.steps, 4), 'higher_is_better': True,
            'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32 (tf32 tensor-core GEMM, fp32 accumulate)',
            'data': 'synthetic',
            'config': bench_config(args.workload, world, nbatch),
            'e2e': {'value': round(e2e_value, 1), 'unit': 'samples/s', 'h2d_bytes_per_step': int(in_bytes),
                    'd2h_bytes_per_step': 4, 'ms_per_step': round(e2e_ms / args.steps, 4), 'timing': e2e_stats,
                    'loss_read': 'every step, one step behind its launch (pinned readback ring, third stream)',
                    'ms_per_step_blocking_read': round(e2e_blocking_ms / args.steps, 4)},
            'gpu_launches': int(launches), 'launches_per_step': round(launches / args.steps, 1),
            'step_graph': graph_info, 'timing': dev_stats,
            'wall_ms_per_step': round(dev_wall / args.steps, 4), 'final_loss': last_loss,
            'clocks': clocks, 'roofline': roofline,
            'kernel_families': dict((k, {'calls_per_step': round(v['calls'] / inst_steps, 1),
                                         'us_per_call': round(1e3 * v['ms'] / v['calls'], 2),
                                         'share': round(v['ms'] / tot, 3),
                                         'replayed_us_per_step': round(1e3 * replayed[k]['ms_per_step'], 2) if k in replayed else None})
                                    for k, v in ranked[:8]),
        }
    return out, rank, world


# ------------------------------------------------------------------ ufunc sweep (config 2) --
def run_ufunc_sweep(sizes=(10 ** 4, 10 ** 5, 10 ** 6, 10 ** 7, 10 ** 8, 10 ** 9)):
    """GB/s of add / multiply / exp / maximum(0,x) / (n/1024,1024)+(1024,) / sum axis 0, 1 / max
    axis 1 / full sum / softmax rows through `minpy.numpy` on one GPU, 1e4 .. 1e9 elements (BASELINE
    config 2).  Every row carries `l2_resident`: true when the bytes the op moves fit the 126 MB L2 --
    such a row is a cache / launch-latency figure, NOT HBM evidence; `hbm_summary` only counts the rest."""
    import minpy
    minpy.set_global_policy('only_cuda')
    import minpy.numpy as mnp
    import minpy.numpy.random as mrandom
    from minpy_b200._lib import get_lib
    lib = get_lib()
    peaks = measured_peaks()
    res = {}
    e0, e1 = ctypes.c_void_p(), ctypes.c_void_p()
    lib.mc_event_create(ctypes.byref(e0))
    lib.mc_event_create(ctypes.byref(e1))

    def time_op(fn, reps):
        for _ in range(3):
            fn()
        lib.mc_sync()
        lib.mc_event_record(e0)
        for _ in range(reps):
            fn()
        lib.mc_event_record(e1)
        lib.mc_sync()
        ms = ctypes.c_float()
        lib.mc_event_elapsed_ms(e0, e1, ctypes.byref(ms))
        return ms.value / reps

    for n in sizes:
        rows = max(1, n // 1024)
        n2 = rows * 1024
        a = mrandom.rand(n2)
        b = mrandom.rand(n2)
        m = a.reshape((rows, 1024))
        v = mrandom.rand(1024)
        reps = 200 if n <= 10 ** 6 else (50 if n <= 10 ** 7 else (10 if n <= 10 ** 8 else 5))
        ops = {
            'add': (lambda: mnp.add(a, b), 12.0 * n2),
            'multiply': (lambda: mnp.multiply(a, b), 12.0 * n2),
            'exp': (lambda: mnp.exp(a), 8.0 * n2),
            'relu': (lambda: mnp.maximum(0, a), 8.0 * n2),
            'bcast_add': (lambda: m + v, 8.0 * n2 + 4096),
            'sum_axis0': (lambda: mnp.sum(m, axis=0), 4.0 * n2 + 4096),
            'sum_axis1': (lambda: mnp.sum(m, axis=1), 4.0 * n2 + 4.0 * rows),
            'max_axis1': (lambda: mnp.max(m, axis=1), 4.0 * n2 + 4.0 * rows),
            'sum_all': (lambda: mnp.sum(a), 4.0 * n2),
        }
        if hasattr(mnp, 'softmax_output'):
            ops['softmax_rows'] = (lambda: mnp.softmax_output(m, m), 8.0 * n2)
        entry = {}
        for name, (fn, nbytes) in ops.items():
            ms = time_op(fn, reps)
            gbs = nbytes / (ms * 1e-3) / 1e9
            entry[name] = {'us': round(ms * 1e3, 2), 'gbs': round(gbs, 1), 'frac_hbm': round(gbs / peaks['hbm_gbs'], 3),
                           'frac_nominal_8tbs': round(gbs / 8000.0, 3), 'l2_resident': bool(nbytes <= L2_BYTES)}
        res['n=%.0e' % n2] = entry
        del a, b, m, v, ops
    beyond = [(k, op, e) for k, ent in res.items() for op, e in ent.items() if not e['l2_resident']]
    if beyond:
        worst = min(beyond, key=lambda t: t[2]['frac_hbm'])
        res['hbm_summary'] = {'rows_beyond_l2': len(beyond),
                              'min_frac_hbm': worst[2]['frac_hbm'], 'min_at': '%s %s' % (worst[0], worst[1]),
                              'median_frac_hbm': round(float(np.median([e['frac_hbm'] for _, _, e in beyond])), 3),
                              'peak_gbs': peaks['hbm_gbs'], 'peak_source': peaks['source']}
    return res


def run_gemm_sweep(sizes=(2048, 4096, 8192)):
    """np.dot (mc_gemm, tcgen05 kind::tf32) at square sizes where the tensor pipe is reachable: TFLOP/s per size and the
    fraction of the kind::tf32 ceiling (half the measured bf16 cuBLAS rate).  Each size alone in a CUDA graph, events;
    operands (>= 50 MB) are re-read from L2 / HBM as the tile raster dictates -- nothing is cached between sizes."""
    from minpy_b200._lib import get_lib
    from minpy_b200.cuda_array import CudaArray
    lib = get_lib()
    peaks = measured_peaks()
    ceiling = peaks['bf16_tflops_sustained'] / 2.0
    e0, e1 = ctypes.c_void_p(), ctypes.c_void_p()
    lib.mc_event_create(ctypes.byref(e0))
    lib.mc_event_create(ctypes.byref(e1))
    rng = np.random.RandomState(0)
    res = {}
    for n in sizes:
        a = rng.randn(n, n).astype(np.float32)
        A, B, C = CudaArray.from_numpy(a), CudaArray.from_numpy(a.T.copy()), CudaArray.empty((n, n))

        def run():
            lib.mc_gemm(0, 0, n, n, n, A.ptr, n, B.ptr, n, C.ptr, n, None, 0, 0, 1)
        for _ in range(3):
            run()
        lib.mc_sync()
        reps = 20 if n <= 4096 else 5
        lib.mc_event_record(e0)
        for _ in range(reps):
            run()
        lib.mc_event_record(e1)
        lib.mc_sync()
        ms = ctypes.c_float()
        lib.mc_event_elapsed_ms(e0, e1, ctypes.byref(ms))
        us = ms.value / reps * 1e3
        rows = rng.randint(0, n, 4)
        ref = a[rows].astype(np.float64) @ a.T.astype(np.float64)
        err = float(np.max(np.abs(C.asnumpy()[rows] - ref)) / np.max(np.abs(ref)))
        tf = 2.0 * n ** 3 / us / 1e6
        res['%d^3' % n] = {'us': round(us, 2), 'tflops': round(tf, 1), 'frac_of_tf32_ceiling': round(tf / ceiling, 3),
                           'relerr_vs_float64': err, 'l2_resident': bool(12.0 * n * n <= L2_BYTES)}
        del A, B, C
    res['tf32_ceiling_tflops'] = round(ceiling, 1)
    res['ceiling_source'] = 'half the %s bf16 cuBLAS sustained rate (MEASURED_PEAKS.json)' % peaks['source']
    return res


# ------------------------------------------------------------- CPU arm (oracle port) -------
def oracle_step_fn(workload, mask_fn=None, with_grads=False, decisions=None):
    """One training step of `workload` on the NumPy oracle (float64, as the reference's NumPy
    variant computes): returns (step(xb, yb) -> loss, batch).  `mask_fn(i, shape)` supplies the
    dropout mask of hidden layer i (parity tests hand in the device's Philox masks); with
    `with_grads` the step returns (loss, {name: gradient}) and `step.params` exposes the state.
    `decisions` = {'relu': [keep masks of layers.relu calls in order], 'conv': [post-relu NCHW outputs
    of the trunk's convolutions]} evaluates the piecewise-linear branch the DEVICE took at every relu /
    max-pool kink (oracle/minpy_oracle.py `relu_decided`); None = the oracle's own decisions."""
    from oracle import minpy_oracle as orc
    from minpy_b200 import workloads
    factory, batch, rule, optim_config, _, _ = workloads.WORKLOADS[workload]
    model = factory()
    params = dict((k, v.astype(np.float64)) for k, v in workloads.init_params(model, 3).items())
    names = list(params.keys())
    cfgs = dict((k, dict(optim_config)) for k in names)
    update = getattr(orc, rule)
    state = {'aux': {}}

    def relu_at(i, a):
        if decisions is None:
            return orc.relu(a)
        return orc.relu_decided(a, np.asarray(decisions['relu'][i], np.float64))

    def conv_block(i, h, w, b, pad):
        a = orc.conv2d(h, w, b, 1, pad)
        if decisions is None:
            return orc.maxpool2d(orc.relu_mx(a), 2, 2)
        dev = np.asarray(decisions['conv'][i], np.float64)      # the device's relu(conv) output, NCHW
        _, arg = orc.maxpool2d_fwd(dev, 2, 2)
        return orc.maxpool2d_decided(orc.relu_decided(a, (dev > 0).astype(np.float64)), arg)

    if workload in ('mlp_c1',):
        def loss_of(xb, yb):
            def f(*ps):
                p = dict(zip(names, ps))
                h = relu_at(0, orc.affine(xb, p['w1'], p['b1']))
                return orc.softmax_loss(orc.affine(h, p['w2'], p['b2']), yb)
            return f
    elif workload == 'mlp_c3':
        depth = model.depth
        rng = np.random.RandomState(17)

        def loss_of(xb, yb):
            def f(*ps):
                p = dict(zip(names, ps))
                h = xb
                for i in range(depth):
                    h = relu_at(i, orc.affine(h, p['w%d' % i], p['b%d' % i]))
                    h, rm, rv = orc.batchnorm(h, p['gamma%d' % i], p['beta%d' % i],
                                              running_mean=state['aux'].get('rm%d' % i),
                                              running_var=state['aux'].get('rv%d' % i))
                    state['aux']['rm%d' % i], state['aux']['rv%d' % i] = orc.raw(rm), orc.raw(rv)
                    shape = orc.raw(h).shape
                    mask = mask_fn(i, shape) if mask_fn is not None else rng.rand(*shape) > 0.5
                    h = orc.dropout(h, 0.5, mask=mask)
                return orc.softmax_loss(orc.affine(h, p['wo'], p['bo']), yb)
            return f
    elif workload in ('cnn_c4', 'cnn_ref'):
        convs = sorted(k[:-7] for k in names if k.endswith('_weight'))
        pad = 1 if workload == 'cnn_c4' else 0

        def loss_of(xb, yb):
            def f(*ps):
                p = dict(zip(names, ps))
                h = xb
                for i, c in enumerate(convs):
                    h = conv_block(i, h, p[c + '_weight'], p[c + '_bias'], pad)
                h = orc.flatten(h)
                if workload == 'cnn_ref':
                    h = relu_at(0, orc.affine(h, p['w1'], p['b1']))
                    return orc.softmax_loss(orc.affine(h, p['w2'], p['b2']), yb)
                return orc.softmax_loss(orc.affine(h, p['w1'], p['b1']), yb)
            return f
    elif workload == 'lstm_c5':
        H = model.H

        def loss_of(xb, yb):
            N, T, _ = xb.shape

            def f(*ps):
                p = dict(zip(names, ps))
                ph, pc = np.zeros((N, H)), np.zeros((N, H))
                hs = orc.zeros((N, 0, H))
                for t in range(T):
                    ph, pc = orc.lstm_step(xb[:, t, :], ph, pc, p['Wx'], p['Wh'], p['b'])
                    hs = orc.append(hs, ph.reshape(N, 1, H), axis=1)
                return orc.temporal_softmax_loss(orc.temporal_affine(hs, p['Wa'], p['ba']), yb, np.ones((N, T)))
            return f
    else:
        raise ValueError(workload)

    def step(xb, yb):
        flat = xb.ndim == 4 and not workload.startswith('cnn')
        xb = xb.reshape(xb.shape[0], -1).astype(np.float64) if flat else xb.astype(np.float64)
        grads, loss = orc.grad_and_loss(loss_of(xb, yb), argnum=range(len(names)))(*[params[n] for n in names])
        for n, g in zip(names, grads):
            params[n], cfgs[n] = update(params[n], g, cfgs[n])
        if with_grads:
            return float(loss), dict((n, np.asarray(orc.raw(g), np.float64)) for n, g in zip(names, grads))
        return float(loss)
    step.params, step.aux, step.configs = params, state['aux'], cfgs
    return step, batch


REF_WORKLOADS = ('mlp_c1', 'mlp_c3', 'lstm_c5')     # the conv trunks need real MXNet on the reference side


def time_reference(workload, steps, warmup, budget_s):
    """samples/s of the UNMODIFIED reference (oracle/_ref, assembled by oracle/build_ref.py) under its
    'only_numpy' policy, in a process of its own (its package is also called `minpy`), on every host
    core.  Returns the cpu_baseline object, or None when oracle/_ref is absent / cannot run `workload`."""
    site = os.path.join(ROOT, 'oracle', '_ref', 'site', 'minpy')
    if workload not in REF_WORKLOADS or not os.path.isdir(site):
        return None
    cmd = [sys.executable, os.path.join(ROOT, 'oracle', 'ref_runner.py'), '--workload', workload, '--steps', str(steps),
           '--warmup', str(warmup), '--budget', str(budget_s)]
    env = dict(os.environ)
    for var in ('OMP_NUM_THREADS', 'MKL_NUM_THREADS', 'OPENBLAS_NUM_THREADS'):   # torchrun pins these to 1
        env.pop(var, None)
    try:
        res = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=600)
        r = json.loads(res.stdout.strip().splitlines()[-1])
    except Exception as exc:  # pylint: disable=broad-except
        log('reference runner failed (%s); falling back to the oracle port' % exc)
        return None
    return {'value': round(r['value'], 1), 'unit': 'samples/s', 'cores': r['cores'], 'kind': 'reference',
            'sample': '%d steps (+%d warm-up) of %s (batch %d) through the UNMODIFIED reference (oracle/_ref: '
                      'minpy.nn.solver.Solver._step under only_numpy, float64), %.1f s' % (
                          r['steps'], r['warmup'], workload, r['batch'], r['seconds']),
            'host_cpus': r['host_cpus'], 'blas': r['blas'], 'numpy': r['numpy'], 'steps': r['steps'],
            'warmup': r['warmup'], 'seconds': r['seconds'], 'batch': r['batch'], 'final_loss': r['final_loss']}


def time_port(workload, budget_s=12.0, max_steps=50, warmup=1):
    """samples/s of the oracle port on this box's host cores over a bounded number of steps."""
    from minpy_b200 import workloads
    from threadpoolctl import threadpool_limits, threadpool_info
    step, batch = oracle_step_fn(workload)
    make_data = workloads.WORKLOADS[workload][4]
    X, Y = make_data(batch * 4, seed=100)
    with threadpool_limits(limits=os.cpu_count()):
        for i in range(warmup):
            step(X[(i % 4) * batch:(i % 4 + 1) * batch], Y[(i % 4) * batch:(i % 4 + 1) * batch])
        t0 = time.perf_counter()
        n = 0
        while n < max_steps:
            j = n % 4
            step(X[j * batch:(j + 1) * batch], Y[j * batch:(j + 1) * batch])
            n += 1
            if time.perf_counter() - t0 > budget_s:
                break
        dt = time.perf_counter() - t0
        threads = max([p.get('num_threads', 1) for p in threadpool_info()] or [1])
    return {'value': round(batch * n / dt, 1), 'unit': 'samples/s', 'cores': int(threads), 'kind': 'port',
            'sample': '%d steps (+%d warm-up) of %s (batch %d) on the NumPy oracle port, float64 as the reference '
                      'NumPy variant, %.1f s' % (n, warmup, workload, batch, dt),
            'host_cpus': os.cpu_count(), 'steps': n, 'warmup': warmup, 'seconds': dt, 'batch': batch}


def time_cpu(workload, budget_s=15.0, max_steps=50, warmup=1):
    """The CPU baseline: the reference itself when oracle/_ref is here, else the oracle port."""
    return time_reference(workload, max_steps, warmup, budget_s) or time_port(workload, budget_s, max_steps, warmup)


def run_reference(args):
    """`--impl reference`: K steps after W warm-up steps (both honoured exactly unless the 120 s budget
    runs out first -- the line's `steps` says what ran) of the reference's CPU path, all host cores."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return None
    cpu = time_cpu(args.workload, budget_s=120.0, max_steps=max(args.steps, 1), warmup=args.warmup)
    batch, n, dt = cpu['batch'], cpu['steps'], cpu['seconds']
    return {
        'impl': 'reference', 'metric': METRIC, 'value': cpu['value'], 'unit': 'samples/s', 'n_gpus': args.gpus,
        'steps': n, 'warmup': cpu['warmup'], 'ms_per_step': round(1e3 * dt / n, 3), 'higher_is_better': True,
        'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
        # the b200 arm's config, key for key (the driver compares them); where this arm ran is in cpu_baseline
        'config': bench_config(args.workload, max(1, args.gpus), args.pool),
        'cpu_baseline': cpu,
        'e2e': {'value': cpu['value'], 'unit': 'samples/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }


def pin_visible_device():
    """One process per GPU: optionally make ONLY this rank's GPU visible to the process before CUDA starts
    (MINPY_PIN_VISIBLE=1; the rank then addresses it as device 0).  Experiment switch for the per-kernel
    slowdown seen in ranks that can see several GPUs."""
    if os.environ.get('MINPY_PIN_VISIBLE') != '1' or int(os.environ.get('WORLD_SIZE', '1')) <= 1:
        return
    local = os.environ.get('LOCAL_RANK', os.environ.get('RANK', '0'))
    visible = os.environ.get('CUDA_VISIBLE_DEVICES')
    ids = visible.split(',') if visible else None
    os.environ['CUDA_VISIBLE_DEVICES'] = ids[int(local)] if ids else local
    os.environ['LOCAL_RANK'] = '0'


def dp_parity(steps=4):
    """N > 1 only, before anything is timed: data-parallel training (one step graph per rank, the
    gradient all-reduce a node of it) of the C1 MLP at its real shape, each rank on its own 256-row
    shard, must equal single-process EAGER training on the concatenated global batch -- every loss in
    nn/layers.py is a batch mean, so the averaged shard gradient IS the full-batch gradient.
    Every rank trains the global-batch run itself (W x 256 rows on one GPU) and compares its own final
    parameters; the worst rank is reported.  `ranks_bitwise_equal`: a checksum of the data-parallel
    parameters agrees on all ranks.  Tolerance 1e-2 (TF32 contractions over different batch splits)."""
    import zlib
    from minpy import array as marray
    from minpy.nn.io import DataBatch
    from minpy.nn.solver import Solver
    from minpy_b200 import parallel, workloads
    from minpy_b200.cuda_array import CudaArray
    world, rank = parallel.world_size(), parallel.rank()
    factory, B, rule, cfg, make_data, _ = workloads.WORKLOADS['mlp_c1']
    cfg = dict(cfg, learning_rate=1e-2)          # large enough for the steps to move the weights
    X, Y = make_data(steps * world * B, seed=7)
    X = X.reshape(steps, world * B, -1)
    Y = Y.reshape(steps, world * B)

    def train(graph, shard):
        model = factory()
        solver = Solver(model, None, None, update_rule=rule, optim_config=dict(cfg), init_rule='gaussian',
                        init_config={'stdvar': 0.01}, verbose=False, step_graph=graph)
        solver.init()
        for k, v in workloads.init_params(model, 3).items():
            model.params[k] = marray.wrap(CudaArray.from_numpy(v))
        for i in range(steps):
            xb, yb = X[i], Y[i]
            if shard:
                xb, yb = xb[rank * B:(rank + 1) * B], yb[rank * B:(rank + 1) * B]
            solver._step(DataBatch([marray.wrap(CudaArray.from_numpy(xb))],  # pylint: disable=protected-access
                                   [marray.wrap(CudaArray.from_numpy(yb))], pad=0, index=i))
        params = dict((k, v.asnumpy()) for k, v in model.params.items())
        return params, dict((k, v) for k, v in workloads.init_params(model, 3).items()), solver

    dp, init, solver = train(True, True)
    sg = getattr(solver, '_step_graph', None)
    saved = dict(parallel._state)  # pylint: disable=protected-access
    parallel._state.update(world=1, native=False)  # pylint: disable=protected-access
    try:
        ref, _, _ = train(False, False)
    finally:
        parallel._state.update(saved)  # pylint: disable=protected-access
    # measured on the update the training made (w_final - w_init): the weights themselves barely move
    err = 0.0
    for k in ref:
        moved = ref[k].astype(np.float64) - init[k]
        diff = dp[k].astype(np.float64) - ref[k]
        err = max(err, float(np.max(np.abs(diff)) / max(1e-30, np.max(np.abs(moved)))))
    crc = 0
    for k in sorted(dp):
        crc = zlib.crc32(dp[k].tobytes(), crc)
    crcs = parallel.allgather_int(crc)
    worst = parallel.max_over_ranks(err)
    return {'max_rel': float('%.3g' % worst), 'tolerance': 1e-2, 'ok': bool(worst <= 1e-2 and len(set(crcs)) == 1),
            'ranks_bitwise_equal': bool(len(set(crcs)) == 1), 'world': world, 'steps': steps,
            'graph_replays': int(sg.replays) if sg is not None else 0,
            'what': 'mlp_c1 (3072-512-10), %d rows per rank, %d steps of sgd_momentum: data-parallel step-graph '
                    'training vs single-process eager training on the global batch; error relative to the '
                    'largest parameter update' % (B, steps)}


EXTRA_WORKLOADS = ('mlp_c1', 'cnn_c4', 'lstm_c5')


def main():
    pin_visible_device()
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=50)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--workload', default='mlp_c3')
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--pool', type=int, default=16, help='distinct batches cycled per rank')
    ap.add_argument('--graph', type=int, default=1, help='replay the step as a CUDA graph (0 = eager launches)')
    ap.add_argument('--prefetch', type=int, default=1, help='e2e leg: NDArrayIter copies batch i+1 on a second stream')
    ap.add_argument('--min-ms', type=float, default=500.0, dest='min_ms',
                    help='repeat the K-step timed window until this much device time is covered (median window reported)')
    ap.add_argument('--max-windows', type=int, default=400, dest='max_windows')
    ap.add_argument('--no-ufunc', action='store_true')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-workloads', action='store_true', help='skip the mlp_c1 / cnn_c4 / lstm_c5 sub-lines')
    ap.add_argument('--no-dp-parity', action='store_true')
    ap.add_argument('--no-family-replay', action='store_true', dest='no_family_replay')
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)

    if args.impl == 'reference':
        out = run_reference(args)
        if out is not None:
            print(json.dumps(out), flush=True)
        return

    from minpy_b200 import parallel
    world = parallel.init()
    dp = None
    if world > 1 and not args.no_dp_parity:
        import minpy
        minpy.set_global_policy('only_cuda')
        dp = dp_parity()
        if parallel.rank() == 0:
            log('dp_parity: %s' % json.dumps(dp))
    out, rank, world = run_b200(args)
    if rank == 0 and dp is not None:
        out['dp_parity'] = dp
    if args.workload == 'mlp_c3' and not args.no_workloads:
        # the other BASELINE configurations, same protocol, so that the driver's records cover the whole
        # metric ("MLP & CNN ... LSTM", at every N): value / ms_per_step / e2e / roofline per workload
        subs = {}
        for name in EXTRA_WORKLOADS:
            sub_args = argparse.Namespace(**vars(args))
            sub_args.workload = name
            sub_args.min_ms = min(args.min_ms, 250.0)
            sub, _, _ = run_b200(sub_args)
            if rank == 0:
                subs[name] = dict((k, sub[k]) for k in ('value', 'unit', 'ms_per_step', 'e2e', 'roofline', 'gpu_launches',
                                                        'launches_per_step', 'final_loss', 'timing', 'step_graph'))
                subs[name]['config'] = sub['config']
        if rank == 0:
            out['workloads'] = subs
    if rank == 0:
        if world == 1 and not args.no_ufunc:
            out['ufunc'] = run_ufunc_sweep()
            out['gemm_tf32'] = run_gemm_sweep()
        if world == 1 and not args.no_cpu_baseline:
            out['cpu_baseline'] = time_cpu(args.workload)
        print(json.dumps(out), flush=True)
    parallel.barrier()
    parallel.shutdown()


if __name__ == '__main__':
    main()
