"""`python tools/host_step.py [workload]`: host-side cost of one CUDA-graph step (what Python does between
two cudaGraphLaunch calls) against the GPU time of the step, plus a cProfile of the replay loop."""
import cProfile, io, os, pstats, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
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
model, solver = bench.build_solver(wl, None, step_graph=True)
bs = [DataBatch([marray.wrap(CudaArray.from_numpy(X[i * batch:(i + 1) * batch]))],
                [marray.wrap(CudaArray.from_numpy(Y[i * batch:(i + 1) * batch]))], pad=0, index=i) for i in range(4)]
lib = get_lib()
for i in range(10):
    solver._step(bs[i % 4])
lib.mc_sync()
n = 300
t0 = time.perf_counter()
for i in range(n):
    solver._step(bs[i % 4])
t1 = time.perf_counter()
lib.mc_sync()
t2 = time.perf_counter()
print('%s: host %.1f us/step, host+drain %.1f us/step (GPU-bound when the second exceeds the first)' % (
    wl, (t1 - t0) / n * 1e6, (t2 - t0) / n * 1e6))
solver.loss_history = []
pr = cProfile.Profile()
pr.enable()
for i in range(n):
    solver._step(bs[i % 4])
pr.disable()
lib.mc_sync()
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats('tottime').print_stats(18)
print(s.getvalue()[:4500])
