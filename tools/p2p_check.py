"""Run under torchrun with W ranks (2..8): the library's two-shot peer-memory all-reduce
(mc_allreduce_avg on a range of the IPC window, csrc/comm.cu) against the exact mean computed on the
host, for sizes that do and do not divide by W, eagerly and as CUDA-graph replays; the result must be
bit-identical on every rank.  Prints 'p2p_check ok' on rank 0."""
import os, sys, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch.distributed as dist
os.environ.setdefault('MINPY_P2P_ALLREDUCE', '1')   # the window at any world size 2..8
import minpy
minpy.set_global_policy('only_cuda')
from minpy_b200 import parallel
from minpy_b200.cuda_array import CudaArray
from minpy_b200._lib import get_lib

world = parallel.init()
rank = parallel.rank()
lib = get_lib()
assert parallel.p2p_active(), 'the peer-memory window is not active'
base, size = parallel._state['p2p']


def contribution(r, n, seed):
    return np.random.RandomState(1000 * seed + r).randn(n).astype(np.float32)


def check(n, off, seed, replay):
    a = CudaArray(parallel._WindowBuffer(base, size), base + 4 * off, (n,))
    mine = contribution(rank, n, seed)
    want = contribution(0, n, seed).copy()
    for q in range(1, world):
        want += contribution(q, n, seed)              # rank order, float32: what the kernel does
    want *= np.float32(1.0 / world)
    if replay:
        src = CudaArray.from_numpy(mine)
        lib.mc_sync()
        lib.mc_graph_begin()
        lib.mc_memcpy_d2d(a.ptr, src.ptr, 4 * n)
        lib.mc_allreduce_avg(a.ptr, n)
        ex = ctypes.c_void_p()
        lib.mc_graph_end(ctypes.byref(ex))
        for _ in range(3):
            lib.mc_graph_launch(ex)
    else:
        lib.mc_memcpy_h2d(a.ptr, mine.ctypes.data, 4 * n)
        lib.mc_allreduce_avg(a.ptr, n)
    got = a.asnumpy()
    assert np.array_equal(got, want), ('mismatch', n, off, replay, float(np.max(np.abs(got - want))))
    gathered = [None] * world
    dist.all_gather_object(gathered, got.tobytes())
    assert all(g == gathered[0] for g in gathered), 'ranks disagree bitwise'


seed = 0
for n, off in ((4, 0), (64, 0), (1024, 64), (113728, 0), (1578496, 128), (2372096, 0), (4 * 1000003, 64)):
    for replay in (False, True):
        seed += 1
        check(n, off, seed, replay)
ptr, nbytes, bad = ctypes.c_void_p(), ctypes.c_size_t(), ctypes.c_int()
lib.mc_p2p_info(ctypes.byref(ptr), ctypes.byref(nbytes), ctypes.byref(bad))
assert bad.value == 0, 'a barrier timed out'
parallel.barrier()
if rank == 0:
    print('p2p_check ok: world %d, window %.0f MB' % (world, size / 2 ** 20), flush=True)
parallel.shutdown()
