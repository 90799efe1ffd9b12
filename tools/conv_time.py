"""Per-layer timing of the C4 convolution trunk (batch 512; 3->32 @32x32, 32->64 @16x16, 64->128 @8x8): every call
alone in a CUDA graph of REPS launches, timed with events.  Prints us/call beside the algorithmic HBM bytes and the
real (un-padded) FLOPs, so the distance to the roofline can be read per layer rather than per kernel family.

    python tools/conv_time.py [batch]
"""
import ctypes
import sys

import numpy as np

from minpy_b200 import ops
from minpy_b200._lib import get_lib
from minpy_b200.cuda_array import CudaArray

REPS = 20


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
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 512
    lib = get_lib()
    rng = np.random.RandomState(0)
    saved = ops.SIDE_STREAM
    ops.SIDE_STREAM = False                     # time the weight-gradient kernels on the main stream
    x = CudaArray.from_numpy(rng.randn(B, 3, 32, 32).astype(np.float32))
    a = ops.to_nhwc(x, 1)
    print('%-34s %9.2f us' % ('nchw_to_nhwc (B,3,32,32)', timed(lib, lambda: ops.to_nhwc(x, 1))))
    for (C, F, HW) in ((3, 32, 32), (32, 64, 16), (64, 128, 8)):
        w = CudaArray.from_numpy((rng.randn(F, C, 3, 3) * 0.05).astype(np.float32))
        b = CudaArray.from_numpy(np.zeros(F, np.float32))
        if a.wpad == 0 and C != 3:
            pass
        y, idx, like = ops.conv2d_nhwc_fwd_pool(a, w, b, 1, True, True)
        flops = 2.0 * B * HW * HW * F * C * 9
        in_b, out_b = B * HW * HW * a.Cp * 4, B * (HW // 2) ** 2 * F * 5
        t = timed(lib, lambda: ops.conv2d_nhwc_fwd_pool(a, w, b, 1, True, True))
        print('%-34s %9.2f us   %6.2f GFLOP  %6.1f MB  -> %5.0f GB/s  %6.1f TFLOP/s'
              % ('fwd+relu+pool %d->%d @%d' % (C, F, HW), t, flops / 1e9, (in_b + out_b) / 1e6, (in_b + out_b) / t / 1e3, flops / t / 1e6))
        dyp = ops.Nhwc(CudaArray.from_numpy(rng.randn(B, HW // 2, HW // 2, F).astype(np.float32)), B, F, HW // 2, HW // 2, F)
        t = timed(lib, lambda: ops.maxpool2d_nhwc_bwd(dyp, idx, like, 2, y))
        mb = (B * (HW // 2) ** 2 * F * 9 + B * HW * HW * F * 4) / 1e6
        print('%-34s %9.2f us   %6.1f MB  -> %5.0f GB/s' % ('  maxpool+relu bwd', t, mb, mb * 1e3 / t))
        dy = ops.maxpool2d_nhwc_bwd(dyp, idx, like, 2, y)
        if C != 3:
            t = timed(lib, lambda: lib.mc_conv2d_nhwc_dgrad(dy.data.ptr, w.ptr, ops.empty((B, HW, HW, a.Cp)).ptr, B, C, HW, HW,
                                                            a.Cp, F, dy.Cp, 3, 3, 1))
            mb = (B * HW * HW * (F + a.Cp) * 4) / 1e6
            print('%-34s %9.2f us   %6.2f GFLOP  %6.1f MB  -> %5.0f GB/s  %6.1f TFLOP/s'
                  % ('  dgrad', t, flops / 1e9, mb, mb * 1e3 / t, flops / t / 1e6))
        dw, db = ops.empty((F, C, 3, 3)), ops.empty((F,))
        t = timed(lib, lambda: lib.mc_conv2d_nhwc_wgrad(a.data.ptr, dy.data.ptr, dw.ptr, db.ptr, B, C, HW, HW, a.Cp, a.wpad,
                                                        F, dy.Cp, 3, 3, 1))
        mb = (B * HW * HW * (F + a.Cp) * 4) / 1e6
        print('%-34s %9.2f us   %6.2f GFLOP  %6.1f MB  -> %5.0f GB/s  %6.1f TFLOP/s'
              % ('  wgrad (+ bias grad)', t, flops / 1e9, mb, mb * 1e3 / t, flops / t / 1e6))
        a = y
    ops.SIDE_STREAM = saved


if __name__ == '__main__':
    main()
