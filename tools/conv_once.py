"""Each C4-shape convolution call twice, eagerly: the target of an ncu launch list
(ncu --metrics gpu__time_duration.sum -k regex:conv python tools/conv_once.py)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from minpy_b200._lib import get_lib
from minpy_b200.cuda_array import CudaArray
lib = get_lib()
N = 512
rng = np.random.RandomState(0)
for name, C, H, F in [('conv1', 3, 32, 32), ('conv2', 32, 16, 64), ('conv3', 64, 8, 128)]:
    Cp = (C + 3) // 4 * 4
    merged = 3 * Cp <= 32
    wpad = 1 if merged else 0
    x = CudaArray.from_numpy(rng.randn(N, H, H + 2 * wpad, Cp).astype(np.float32))
    w = CudaArray.from_numpy(rng.randn(F, C, 3, 3).astype(np.float32))
    b = CudaArray.from_numpy(np.zeros(F, np.float32))
    y = CudaArray.empty((N, H, H, F)); dx = CudaArray.empty((N, H, H, Cp)); dw = CudaArray.empty((F, C, 3, 3)); db = CudaArray.empty((F,))
    for _ in range(2):
        lib.mc_conv2d_nhwc_fwd(x.ptr, w.ptr, b.ptr, y.ptr, N, C, H, H, Cp, wpad, F, F, 3, 3, 1, 1)
        if not merged:
            lib.mc_conv2d_nhwc_dgrad(y.ptr, w.ptr, dx.ptr, N, C, H, H, Cp, F, F, 3, 3, 1)
        lib.mc_conv2d_nhwc_wgrad(x.ptr, y.ptr, dw.ptr, db.ptr, N, C, H, H, Cp, wpad, F, F, 3, 3, 1)
    lib.mc_sync()
