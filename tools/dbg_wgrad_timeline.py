"""MC_CONV_DBG=1 python tools/dbg_wgrad_timeline.py: clock64 timeline of CTA 0 of the halo weight-gradient kernel."""
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
    wpad = 1 if 3 * Cp <= 32 else 0
    x = CudaArray.from_numpy(rng.randn(N, H, H + 2 * wpad, Cp).astype(np.float32))
    y = CudaArray.from_numpy(rng.randn(N, H, H, F).astype(np.float32))
    dw = CudaArray.empty((F, C, 3, 3)); db = CudaArray.empty((F,))
    for i in range(2):
        sys.stderr.write('== %s wgrad run %d\n' % (name, i))
        lib.mc_conv2d_nhwc_wgrad(x.ptr, y.ptr, dw.ptr, db.ptr, N, C, H, H, Cp, wpad, F, F, 3, 3, 1)
        lib.mc_sync()
