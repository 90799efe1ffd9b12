"""mc_conv2d_nhwc_{fwd,dgrad,wgrad} timed through the C ABI at the C4 layer shapes (B=512)."""
import sys, ctypes, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from minpy_b200._lib import get_lib
from minpy_b200.cuda_array import CudaArray
lib = get_lib()
e0, e1 = ctypes.c_void_p(), ctypes.c_void_p()
lib.mc_event_create(ctypes.byref(e0)); lib.mc_event_create(ctypes.byref(e1))
def timeit(fn, reps=20, inner=10):
    """GPU time per call: `inner` calls captured into a CUDA graph (no host gaps), `reps` launches."""
    for _ in range(3): fn()
    lib.mc_sync()
    lib.mc_graph_begin()
    for _ in range(inner): fn()
    ex = ctypes.c_void_p(); lib.mc_graph_end(ctypes.byref(ex))
    lib.mc_graph_launch(ex)
    lib.mc_sync(); lib.mc_event_record(e0)
    for _ in range(reps): lib.mc_graph_launch(ex)
    lib.mc_event_record(e1); lib.mc_sync()
    ms = ctypes.c_float(); lib.mc_event_elapsed_ms(e0, e1, ctypes.byref(ms))
    lib.mc_graph_destroy(ex)
    return ms.value / (reps * inner) * 1e3
N = 512
rng = np.random.RandomState(0)
for name, C, H, F in [('conv1', 3, 32, 32), ('conv2', 32, 16, 64), ('conv3', 64, 8, 128)]:
    Cp, Fp = (C + 3) // 4 * 4, F
    merged = 3 * Cp <= 32
    wpad = 1 if merged else 0
    x = CudaArray.from_numpy(rng.randn(N, H, H + 2 * wpad, Cp).astype(np.float32))
    w = CudaArray.from_numpy(rng.randn(F, C, 3, 3).astype(np.float32))
    b = CudaArray.from_numpy(np.zeros(F, np.float32))
    y = CudaArray.empty((N, H, H, Fp)); dx = CudaArray.empty((N, H, H, Cp)); dw = CudaArray.empty((F, C, 3, 3)); db = CudaArray.empty((F,))
    flops = 2.0 * N * H * H * F * C * 9
    nbytes = 4.0 * (N * H * H * Cp + N * H * H * Fp)
    t = timeit(lambda: lib.mc_conv2d_nhwc_fwd(x.ptr, w.ptr, b.ptr, y.ptr, N, C, H, H, Cp, wpad, F, Fp, 3, 3, 1, 1))
    print('%s fwd   %8.2f us  %6.1f TF/s  %7.1f GB/s' % (name, t, flops / t / 1e6, nbytes / t / 1e3), flush=True)
    if not merged:
        t = timeit(lambda: lib.mc_conv2d_nhwc_dgrad(y.ptr, w.ptr, dx.ptr, N, C, H, H, Cp, F, Fp, 3, 3, 1))
        print('%s dgrad %8.2f us  %6.1f TF/s  %7.1f GB/s' % (name, t, flops / t / 1e6, nbytes / t / 1e3), flush=True)
    t = timeit(lambda: lib.mc_conv2d_nhwc_wgrad(x.ptr, y.ptr, dw.ptr, db.ptr, N, C, H, H, Cp, wpad, F, Fp, 3, 3, 1))
    print('%s wgrad %8.2f us  %6.1f TF/s  %7.1f GB/s' % (name, t, flops / t / 1e6, nbytes / t / 1e3), flush=True)

# pooling and layout conversion at the same layer shapes
for name, Cp, H in [('pool1', 32, 32), ('pool2', 64, 16), ('pool3', 128, 8)]:
    x = CudaArray.from_numpy(rng.randn(N, H, H, Cp).astype(np.float32))
    y = CudaArray.empty((N, H // 2, H // 2, Cp)); dx = CudaArray.empty((N, H, H, Cp))
    idx = CudaArray.empty((N * (H // 2) * (H // 2) * Cp // 4,))
    nb_f = 4.0 * N * H * H * Cp * 1.25 + N * H * H * Cp / 4
    t = timeit(lambda: lib.mc_maxpool2d_nhwc_fwd(x.ptr, y.ptr, idx.ptr, N, H, H, Cp, 2, 2, 2))
    print('%s fwd   %8.2f us  %7.1f GB/s' % (name, t, nb_f / t / 1e3), flush=True)
    nb_b = 4.0 * N * H * H * Cp * 1.5 + N * H * H * Cp / 4
    t = timeit(lambda: lib.mc_maxpool2d_nhwc_bwd(y.ptr, idx.ptr, y.ptr, dx.ptr, N, H, H, Cp, 2, H // 2, H // 2))
    print('%s bwd   %8.2f us  %7.1f GB/s' % (name, t, nb_b / t / 1e3), flush=True)
xin = CudaArray.from_numpy(rng.randn(N, 3, 32, 32).astype(np.float32)); xo = CudaArray.empty((N, 32, 34, 4))
t = timeit(lambda: lib.mc_nchw_to_nhwc(xin.ptr, xo.ptr, N, 3, 32, 32, 4, 1))
print('input nchw->nhwc4 %8.2f us  %7.1f GB/s' % (t, 4.0 * N * 32 * (32 * 3 + 34 * 4) / t / 1e3), flush=True)
a = CudaArray.from_numpy(rng.randn(N, 4, 4, 128).astype(np.float32)); bb = CudaArray.empty((N, 128, 4, 4))
t = timeit(lambda: lib.mc_nhwc_to_nchw(a.ptr, bb.ptr, N, 128, 4, 4, 128))
print('trunk nhwc->nchw  %8.2f us  %7.1f GB/s' % (t, 8.0 * N * 2048 / t / 1e3), flush=True)
t = timeit(lambda: lib.mc_nchw_to_nhwc(bb.ptr, a.ptr, N, 128, 4, 4, 128, 0))
print('trunk nchw->nhwc  %8.2f us  %7.1f GB/s' % (t, 8.0 * N * 2048 / t / 1e3), flush=True)
