import sys, numpy as np
sys.path.insert(0, '.')
from minpy_b200 import ops
from minpy_b200.cuda_array import CudaArray
from oracle import minpy_oracle as orc
N, C, H, W, F, K, pad = [int(v) for v in sys.argv[1:8]]
rng = np.random.RandomState(0)
x = rng.randn(N, C, H, W).astype(np.float32)
w = (rng.randn(F, C, K, K) / np.sqrt(C * K * K)).astype(np.float32)
y = ops.conv2d_fwd(CudaArray.from_numpy(x), CudaArray.from_numpy(w), None, 1, pad).asnumpy()
ref = orc.conv2d_fwd(x.astype(np.float64), w.astype(np.float64), None, 1, pad)
print(sys.argv[1:8], 'rel err', np.max(np.abs(y - ref)) / np.max(np.abs(ref)))
