"""Phase timing of head_softmax_kernel from inside the kernel (library built by tools/head_probe.sh).

    MINPY_CUDA_LIB=tools/_probe/libminpy_cuda_probe.so python tools/head_probe.py [rows K C]

Stamps (thread 0 of each block, %globaltimer ns): 0 entry, 1 after the PDL wait, 2 W in shared memory,
3 logits / softmax done, 4 ticket taken and dh written (thread 0), 5 = 6 after the closing barrier,
7 last block's sum stored.
"""
import ctypes
import sys

import numpy as np

from minpy_b200 import ops
from minpy_b200._lib import get_lib
from minpy_b200.cuda_array import CudaArray


def main():
    rows, K, C = (int(v) for v in sys.argv[1:4]) if len(sys.argv) >= 4 else (1024, 512, 10)
    lib = get_lib()
    rng = np.random.RandomState(0)
    h = CudaArray.from_numpy(rng.randn(rows, K).astype(np.float32))
    W = CudaArray.from_numpy((rng.randn(K, C) * 0.05).astype(np.float32))
    b = CudaArray.from_numpy(np.zeros(C, np.float32))
    y = CudaArray.from_numpy(rng.randint(0, C, rows).astype(np.int32))
    logits, dl, dh, loss = ops.empty((rows, C)), ops.empty((rows, C)), ops.empty((rows, K)), ops.empty(())
    read = lib.cdll.mc_head_probe_read
    read.argtypes = [ctypes.c_void_p]
    buf = np.zeros(1024 * 8, np.uint64)
    blocks = min(1024, (rows + 7) // 8)   # HEAD_ROWS = 8
    for it in range(6):
        lib.mc_head_softmax_loss(h.ptr, W.ptr, b.ptr, y.ptr, logits.ptr, dl.ptr, dh.ptr, loss.ptr, rows, K, C)
        lib.mc_sync()
        assert read(buf.ctypes.data) == 0
        t = buf.reshape(1024, 8)[:blocks].astype(np.int64)
        t0 = t[:, 0].min()
        rel = t - t0
        ph = np.diff(t[:, :7], axis=1)
        print('it %d: first block entry 0, last block entry %+d ns; block end (stamp 6) min %d max %d ns; '
              'loss stored at %d ns' % (it, rel[:, 0].max(), rel[:, 6].min(), rel[:, 6].max(), t[:, 7].max() - t0))
        print('   per-phase ns, median over blocks [pdl, W+h load, logits+softmax, ticket+dh, barrier, -]: %s   max: %s'
              % (np.median(ph, axis=0).astype(int).tolist(), ph.max(axis=0).tolist()))


if __name__ == '__main__':
    main()
