import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
def P(*a): print('[r%s %.2f]' % (os.environ.get('RANK'), time.time() % 1000), *a, file=sys.stderr, flush=True)
import minpy
minpy.set_global_policy('only_cuda')
from minpy_b200 import parallel
from minpy_b200.cuda_array import CudaArray
from minpy_b200._lib import get_lib
import numpy as np
P('init...')
world = parallel.init()
P('init done native', parallel.native_comm())
lib = get_lib()
a = CudaArray.from_numpy(np.full(1024, parallel.rank() + 1.0, np.float32))
lib.mc_allreduce_avg(a.ptr, a.size)
P('allreduce enqueued')
lib.mc_sync()
P('allreduce value', a.asnumpy()[:2])
parallel.barrier()
P('barrier ok')
lib.mc_graph_begin()
lib.mc_allreduce_avg(a.ptr, a.size)
import ctypes
ex = ctypes.c_void_p()
lib.mc_graph_end(ctypes.byref(ex))
P('captured')
for i in range(3):
    lib.mc_graph_launch(ex)
lib.mc_sync()
P('graph allreduce value', a.asnumpy()[:2])
parallel.barrier()
parallel.shutdown()
P('done')
