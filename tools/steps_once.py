"""`python tools/steps_once.py <workload> [steps]`: a few Solver steps of one bench workload (2 eager,
then CUDA-graph replays) with device-resident batches -- the short target for ncu launch lists and
full captures (bench.py itself adds an eager leg, a roofline replay and the CPU baseline)."""
import sys, os
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
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 6
_, batch, _, _, make_data, _ = workloads.WORKLOADS[wl]
X, Y = make_data(batch * 2, seed=1)
model, solver = bench.build_solver(wl, None, step_graph=True)
bs = [DataBatch([marray.wrap(CudaArray.from_numpy(X[i * batch:(i + 1) * batch]))],
                [marray.wrap(CudaArray.from_numpy(Y[i * batch:(i + 1) * batch]))], pad=0, index=i) for i in range(2)]
for i in range(steps):
    solver._step(bs[i % 2])
get_lib().mc_sync()
print('done', wl, steps, float(solver.loss_history[-1]))
