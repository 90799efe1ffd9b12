"""Deferred loss read (mc_readback_begin / mc_readback_wait, CudaArray.begin_readback): the value that
arrives through the pinned ring is bit-identical to a blocking read, for eager steps and graph
replays, and a ticket that fell out of the ring falls back to the blocking read."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_readback_ring_values_and_stale_ticket():
    from minpy_b200._lib import get_lib
    from minpy_b200.cuda_array import CudaArray
    lib = get_lib()
    vals = np.arange(300, dtype=np.float32) * 0.5 - 7.0
    arrs = [CudaArray.from_numpy(np.float32(v)).begin_readback() for v in vals]
    # the first 44 tickets have been overwritten by the time all 300 copies were started
    out = np.empty((), np.float32)
    assert lib.mc_readback_wait.raw(arrs[0]._readback, out.ctypes.data, 4) == 4
    assert lib.mc_readback_wait.raw(arrs[299]._readback, out.ctypes.data, 4) == 0 and out == vals[299]
    got = np.array([float(a) for a in arrs], np.float32)       # stale ones read the device value
    assert np.array_equal(got, vals)
    assert arrs[0]._readback is None and arrs[299]._readback is not None
    # a vector of up to 64 bytes travels whole; larger values are left to the blocking path
    v16 = CudaArray.from_numpy(np.arange(16, dtype=np.float32)).begin_readback()
    assert v16._readback is not None and np.array_equal(v16.asnumpy(), np.arange(16, dtype=np.float32))
    big = CudaArray.from_numpy(np.arange(17, dtype=np.float32)).begin_readback()
    assert big._readback is None and np.array_equal(big.asnumpy(), np.arange(17, dtype=np.float32))
    # a ticket that was never issued is refused, not waited for
    assert lib.mc_readback_wait.raw(ctypes.c_uint64(1 << 40), out.ctypes.data, 4) == 4


@pytest.mark.parametrize('graph', [False, True])
def test_solver_losses_arrive_through_the_ring(graph):
    """Every entry of loss_history carries a ticket; reading entry i after step i+1 was enqueued gives
    the same number as a blocking read of the device scalar."""
    import minpy
    minpy.set_global_policy('only_cuda')
    from minpy import array as marray
    from minpy.nn.io import NDArrayIter
    from minpy.nn.solver import Solver
    from minpy_b200 import workloads
    from minpy_b200.cuda_array import CudaArray
    factory, _, rule, optim_config, make_data, _ = workloads.WORKLOADS['mlp_c1'][:6]
    X, Y = make_data(64 * 6, seed=3)
    it = NDArrayIter(X, Y, batch_size=64, shuffle=False, num_parts=1, part_index=0)
    solver = Solver(factory(), it, it, update_rule=rule, optim_config=dict(optim_config), init_rule='gaussian',
                    init_config={'stdvar': 0.01}, verbose=False, step_graph=graph)
    solver.init()
    lagged = []
    for _epoch in range(2):
        it.reset()
        for batch in it:
            solver._step(batch)                                  # pylint: disable=protected-access
            if len(solver.loss_history) >= 2:
                lagged.append(float(solver.loss_history[-2]))    # one step behind: waits for that step only
    lagged.append(float(solver.loss_history[-1]))
    assert len(lagged) == len(solver.loss_history) == 12
    for entry, value in zip(solver.loss_history, lagged):
        raw = entry._data[marray.ArrayType.CUDA]                 # pylint: disable=protected-access
        assert isinstance(raw, CudaArray)
        blocking = CudaArray(raw.buf, raw.ptr, raw.shape, raw.dtype).asnumpy()    # a view without a ticket
        assert np.float32(value) == blocking.reshape(-1)[0]
    assert np.isfinite(lagged).all() and len(set(lagged)) > 1
