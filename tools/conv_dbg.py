import numpy as np, sys
from minpy_b200 import ops
from minpy_b200._lib import get_lib
from minpy_b200.cuda_array import CudaArray
lib = get_lib(); rng = np.random.RandomState(0); B = 512
ops.SIDE_STREAM = False
x = CudaArray.from_numpy(rng.randn(B, 3, 32, 32).astype(np.float32))
a = ops.to_nhwc(x, 1)
for (C, F, HW) in ((3, 32, 32), (32, 64, 16), (64, 128, 8)):
    w = CudaArray.from_numpy((rng.randn(F, C, 3, 3) * 0.05).astype(np.float32))
    b = CudaArray.from_numpy(np.zeros(F, np.float32))
    sys.stderr.write('=== fwd+pool %d->%d @%d\n' % (C, F, HW)); sys.stderr.flush()
    for _ in range(2):
        y, idx, like = ops.conv2d_nhwc_fwd_pool(a, w, b, 1, True, True)
    lib.mc_sync()
    dyp = ops.Nhwc(CudaArray.from_numpy(rng.randn(B, HW // 2, HW // 2, F).astype(np.float32)), B, F, HW // 2, HW // 2, F)
    dy = ops.maxpool2d_nhwc_bwd(dyp, idx, like, 2, y)
    if C != 3:
        sys.stderr.write('=== dgrad\n'); sys.stderr.flush()
        lib.mc_conv2d_nhwc_dgrad(dy.data.ptr, w.ptr, ops.empty((B, HW, HW, a.Cp)).ptr, B, C, HW, HW, a.Cp, F, dy.Cp, 3, 3, 1)
    dw, db = ops.empty((F, C, 3, 3)), ops.empty((F,))
    sys.stderr.write('=== wgrad\n'); sys.stderr.flush()
    lib.mc_conv2d_nhwc_wgrad(a.data.ptr, dy.data.ptr, dw.ptr, db.ptr, B, C, HW, HW, a.Cp, a.wpad, F, dy.Cp, 3, 3, 1)
    lib.mc_sync()
    a = y
