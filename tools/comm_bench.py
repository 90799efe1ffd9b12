"""Time mc_allreduce_avg (library NCCL communicator) eagerly and as a CUDA-graph node. torchrun."""
import os, sys, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
os.environ.setdefault('MINPY_P2P_ALLREDUCE', '1')   # time the peer-memory window at any world size
import minpy
minpy.set_global_policy('only_cuda')
from minpy_b200 import parallel
from minpy_b200.cuda_array import CudaArray
from minpy_b200._lib import get_lib
world = parallel.init()
lib = get_lib()
e0, e1 = ctypes.c_void_p(), ctypes.c_void_p()
lib.mc_event_create(ctypes.byref(e0)); lib.mc_event_create(ctypes.byref(e1))
def elapsed(fn, reps):
    lib.mc_sync(); parallel.barrier(); lib.mc_event_record(e0)
    for _ in range(reps): fn()
    lib.mc_event_record(e1); lib.mc_sync()
    ms = ctypes.c_float(); lib.mc_event_elapsed_ms(e0, e1, ctypes.byref(ms))
    return ms.value / reps * 1e3
WINDOW = parallel._state.get('p2p')   # peer-memory window (two-shot all-reduce) or None
cases = [(n, False) for n in (113738, 1578506, 2372106, 4853888)]
if WINDOW is not None:
    cases += [(n, True) for n in (113728, 1578496, 2372096, 4853888)]
for n, in_window in cases:
    if in_window:
        a = CudaArray(parallel._WindowBuffer(*WINDOW), WINDOW[0], (n,))
        lib.mc_fill(a.ptr, 1.0, n)
    else:
        a = CudaArray.from_numpy(np.ones(n, np.float32))
    for _ in range(5): lib.mc_allreduce_avg(a.ptr, n)
    eager = elapsed(lambda: lib.mc_allreduce_avg(a.ptr, n), 50)
    lib.mc_sync()
    lib.mc_graph_begin()
    for _ in range(10): lib.mc_allreduce_avg(a.ptr, n)
    ex = ctypes.c_void_p(); lib.mc_graph_end(ctypes.byref(ex))
    lib.mc_graph_launch(ex)
    graph = elapsed(lambda: lib.mc_graph_launch(ex), 10) / 10
    if parallel.rank() == 0:
        print('world %d  %-4s n %8d (%.2f MB)  eager %7.1f us  in-graph %7.1f us  algbw %.0f GB/s' % (
            world, 'p2p' if in_window else 'nccl', n, 4 * n / 1e6, eager, graph, 4 * n / graph / 1e3), flush=True)
parallel.barrier()
parallel.shutdown()
