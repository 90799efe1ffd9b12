"""Kernel timeline of ONE replayed training step (torch.profiler / CUPTI sees the library's kernels: same process):
start, duration, stream and the gap to the previous kernel on the critical path -- where a step's time goes when
the per-family sums (bench.py) exceed or undershoot it because of the side stream.

    python tools/timeline.py [mlp_c3|mlp_c1|cnn_c4|lstm_c5] [max rows]
"""
import sys

sys.path.insert(0, '.')
import minpy
minpy.set_global_policy('only_cuda')
from minpy import array as marray
from minpy.nn.io import DataBatch
from minpy_b200 import workloads
from minpy_b200.cuda_array import CudaArray
from minpy_b200._lib import get_lib
import bench


def main():
    wl = sys.argv[1] if len(sys.argv) > 1 else 'mlp_c3'
    limit = int(sys.argv[2]) if len(sys.argv) > 2 else 80
    import torch
    from torch.profiler import profile, ProfilerActivity
    _, batch, _, _, make_data, _ = workloads.WORKLOADS[wl]
    X, Y = make_data(batch * 4, seed=1)
    _, solver = bench.build_solver(wl, None, step_graph=True)
    bs = [DataBatch([marray.wrap(CudaArray.from_numpy(X[i * batch:(i + 1) * batch]))],
                    [marray.wrap(CudaArray.from_numpy(Y[i * batch:(i + 1) * batch]))], pad=0, index=i) for i in range(4)]
    lib = get_lib()
    for i in range(12):
        solver._step(bs[i % 4])          # pylint: disable=protected-access
    lib.mc_sync()
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for i in range(6):
            solver._step(bs[i % 4])      # pylint: disable=protected-access
        lib.mc_sync()
        torch.cuda.synchronize()
    ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and e.device_time > 0]
    ev.sort(key=lambda e: e.time_range.start)
    if not ev:
        print('no CUDA events recorded')
        return
    # split into steps at the graph's first kernel (the epoch counter increment)
    starts = [i for i, e in enumerate(ev) if 'graph_epoch_inc' in e.name]
    if len(starts) >= 3:
        a, b = starts[-2], starts[-1]
    else:
        a, b = 0, len(ev)
    step = ev[a:b]
    t0 = step[0].time_range.start
    print('%s: %d kernels / copies in the step, %.1f us from first start to last end' %
          (wl, len(step), max(e.time_range.end for e in step) - t0))
    print('%9s %8s %7s  %s' % ('start us', 'dur us', 'gap us', 'kernel'))
    last_end = t0
    for e in step[:limit]:
        s, d = e.time_range.start - t0, e.time_range.end - e.time_range.start
        print('%9.1f %8.1f %7.1f  %s' % (s, d, e.time_range.start - last_end, e.name[:90]))
        last_end = max(last_end, e.time_range.end)


if __name__ == '__main__':
    main()
