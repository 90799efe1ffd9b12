"""Time mc_lstm_seq_fwd / mc_lstm_seq_bwd alone at the C5 shape (T=100, N=128, H=1024; examples/nn/lstm.py), each in a CUDA graph
of REPS calls, with events.  MC_LSTM_PERSIST=0 gives the launch-per-step path.

    python tools/lstm_time.py [T N H]
"""
import ctypes
import sys

import numpy as np

from minpy_b200._lib import get_lib
from minpy_b200 import ops
from minpy_b200.cuda_array import CudaArray

REPS = 4


def timed(lib, fn):
    fn()
    lib.mc_sync()
    lib.mc_graph_begin()
    for _ in range(REPS):
        fn()
    ex = ctypes.c_void_p()
    lib.mc_graph_end(ctypes.byref(ex))
    for _ in range(2):
        lib.mc_graph_launch(ex)
    e0, e1 = ctypes.c_void_p(), ctypes.c_void_p()
    lib.mc_event_create(ctypes.byref(e0))
    lib.mc_event_create(ctypes.byref(e1))
    lib.mc_sync()
    lib.mc_event_record(e0)
    for _ in range(5):
        lib.mc_graph_launch(ex)
    lib.mc_event_record(e1)
    lib.mc_sync()
    ms = ctypes.c_float()
    lib.mc_event_elapsed_ms(e0, e1, ctypes.byref(ms))
    lib.mc_graph_destroy(ex)
    return ms.value * 1e3 / (5 * REPS)


def main():
    T, N, H = (int(v) for v in sys.argv[1:4]) if len(sys.argv) >= 4 else (100, 128, 1024)
    lib = ops.L()
    rng = np.random.RandomState(0)
    A = CudaArray.from_numpy(rng.randn(T, N, 4 * H).astype(np.float32))
    Wh = CudaArray.from_numpy((rng.randn(H, 4 * H) / np.sqrt(H)).astype(np.float32))
    hb = CudaArray.from_numpy(np.zeros((T + 1, N, H), np.float32))
    cb = CudaArray.from_numpy(np.zeros((T + 1, N, H), np.float32))
    gt = CudaArray.from_numpy(np.zeros((T, N, 4 * H), np.float32))
    dH = CudaArray.from_numpy((rng.randn(T + 1, N, H) * 0.1).astype(np.float32))
    dA = CudaArray.from_numpy(np.zeros((T, N, 4 * H), np.float32))
    prec = lib.MC_GEMM_TF32
    us = timed(lib, lambda: lib.mc_lstm_seq_fwd(A.ptr, Wh.ptr, hb.ptr, cb.ptr, gt.ptr, T, N, H, prec))
    print('mc_lstm_seq_fwd T=%d N=%d H=%d: %9.1f us/call  %6.2f us/step' % (T, N, H, us, us / T))
    us = timed(lib, lambda: lib.mc_lstm_seq_bwd(gt.ptr, cb.ptr, Wh.ptr, dH.ptr, dA.ptr, T, N, H, prec))
    print('mc_lstm_seq_bwd T=%d N=%d H=%d: %9.1f us/call  %6.2f us/step' % (T, N, H, us, us / T))


if __name__ == '__main__':
    main()
