"""cProfile of Solver._step on the cuda variant (host-side overhead hunt)."""
import cProfile, pstats, sys, io
sys.path.insert(0, '.')
import minpy
minpy.set_global_policy('only_cuda')
from minpy import array as marray
from minpy.nn.io import DataBatch
from minpy_b200 import workloads
from minpy_b200.cuda_array import CudaArray
from minpy_b200._lib import get_lib
import bench
wl = sys.argv[1] if len(sys.argv) > 1 else 'mlp_c3'
_, batch, _, _, make_data, _ = workloads.WORKLOADS[wl]
X, Y = make_data(batch * 4, seed=1)
model, solver = bench.build_solver(wl, None)
bs = [DataBatch([marray.wrap(CudaArray.from_numpy(X[i*batch:(i+1)*batch]))], [marray.wrap(CudaArray.from_numpy(Y[i*batch:(i+1)*batch]))], pad=0, index=i) for i in range(4)]
lib = get_lib()
for i in range(10):
    solver._step(bs[i % 4])
lib.mc_sync()
import time
t0 = time.perf_counter()
for i in range(100):
    solver._step(bs[i % 4])
t1 = time.perf_counter()
lib.mc_sync()
t2 = time.perf_counter()
print('host time per step %.3f ms; + drain %.3f ms' % ((t1 - t0) * 10, (t2 - t1) * 1e3))
pr = cProfile.Profile()
pr.enable()
for i in range(100):
    solver._step(bs[i % 4])
pr.disable()
lib.mc_sync()
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats('tottime').print_stats(28)
print(s.getvalue()[:6000])
