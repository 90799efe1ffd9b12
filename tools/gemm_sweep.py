"""mc_gemm (tcgen05 kind::tf32) timed through the C ABI at the shapes of SURVEY 8a (a8) and at
square sizes where the tensor pipe can be reached.  CUDA events on the library stream."""
import sys, ctypes, json
sys.path.insert(0, '.')
import numpy as np
from minpy_b200._lib import get_lib
from minpy_b200.cuda_array import CudaArray

lib = get_lib()
e0, e1 = ctypes.c_void_p(), ctypes.c_void_p()
lib.mc_event_create(ctypes.byref(e0))
lib.mc_event_create(ctypes.byref(e1))

# (name, ta, tb, M, N, K)
SHAPES = [
    ('c1_fwd1', 0, 0, 256, 512, 3072), ('c1_dw1', 1, 0, 3072, 512, 256),
    ('c3_fwd1', 0, 0, 1024, 512, 3072), ('c3_dw1', 1, 0, 3072, 512, 1024),
    ('c3_fwd2', 0, 0, 1024, 512, 512), ('c3_dx2', 0, 1, 1024, 512, 512), ('c3_dw2', 1, 0, 512, 512, 1024),
    ('c5_hWh', 0, 0, 128, 4096, 1024), ('c5_dh', 0, 1, 128, 1024, 4096), ('c5_dWh', 1, 0, 1024, 4096, 128),
    ('c5_vocab', 0, 0, 12800, 128, 1024),
    ('c5_xWx', 0, 0, 12800, 4096, 128), ('c5_dhs', 0, 1, 12800, 1024, 128), ('c5_dWh_all', 1, 0, 1024, 4096, 12800),
    ('c5_dWx', 1, 0, 128, 4096, 12800), ('c5_dx', 0, 1, 12800, 128, 4096), ('c5_dWa', 1, 0, 1024, 128, 12800),
    ('c3_out', 0, 0, 1024, 10, 512), ('c3_dwo', 1, 0, 512, 10, 1024), ('c3_dho', 0, 1, 1024, 512, 10),
    ('c4_out', 0, 0, 512, 10, 2048), ('c4_dwo', 1, 0, 2048, 10, 512), ('c4_dho', 0, 1, 512, 2048, 10),
    ('sq1024', 0, 0, 1024, 1024, 1024), ('sq2048', 0, 0, 2048, 2048, 2048),
    ('sq4096', 0, 0, 4096, 4096, 4096), ('sq4096_nt', 0, 1, 4096, 4096, 4096),
    ('sq4096_tn', 1, 0, 4096, 4096, 4096), ('sq8192', 0, 0, 8192, 8192, 8192),
]
only = sys.argv[1:]
rng = np.random.RandomState(0)
out = {}
for name, ta, tb, M, N, K in SHAPES:
    if only and name not in only:
        continue
    a_shape = (K, M) if ta else (M, K)
    b_shape = (N, K) if tb else (K, N)
    A = CudaArray.from_numpy(rng.randn(*a_shape).astype(np.float32))
    B = CudaArray.from_numpy(rng.randn(*b_shape).astype(np.float32))
    C = CudaArray.from_numpy(np.zeros((M, N), np.float32))

    def run():
        lib.mc_gemm(ta, tb, M, N, K, A.ptr, a_shape[1], B.ptr, b_shape[1], C.ptr, N, None, 0, 0, 1)
    for _ in range(3):
        run()
    lib.mc_sync()
    reps = 20 if M * N * K < 2 ** 34 else 5
    inner = 10 if M * N * K < 2 ** 32 else 1
    lib.mc_graph_begin()                      # `inner` calls per graph: GPU time without host gaps
    for _ in range(inner):
        run()
    ex = ctypes.c_void_p()
    lib.mc_graph_end(ctypes.byref(ex))
    lib.mc_graph_launch(ex)
    lib.mc_sync()
    lib.mc_event_record(e0)
    for _ in range(reps):
        lib.mc_graph_launch(ex)
    lib.mc_event_record(e1)
    lib.mc_sync()
    ms = ctypes.c_float()
    lib.mc_event_elapsed_ms(e0, e1, ctypes.byref(ms))
    lib.mc_graph_destroy(ex)
    us = ms.value / (reps * inner) * 1e3
    flops = 2.0 * M * N * K
    nbytes = 4.0 * (M * K + K * N + M * N)
    # spot check against float64 on a few entries
    got = C.asnumpy()
    an = A.asnumpy().astype(np.float64); bn = B.asnumpy().astype(np.float64)
    if ta: an = an.T
    if tb: bn = bn.T
    rows = rng.randint(0, M, 4)
    ref = an[rows] @ bn
    err = float(np.max(np.abs(got[rows] - ref)) / np.max(np.abs(ref)))
    out[name] = dict(us=round(us, 2), tflops=round(flops / us / 1e6, 1), gbs=round(nbytes / us / 1e3, 1), relerr=err)
    print('%-10s ta=%d tb=%d M=%5d N=%5d K=%5d  %9.2f us  %7.1f TF/s  %7.1f GB/s  err %.1e' %
          (name, ta, tb, M, N, K, us, flops / us / 1e6, nbytes / us / 1e3, err), flush=True)
    del A, B, C
print(json.dumps(out))
