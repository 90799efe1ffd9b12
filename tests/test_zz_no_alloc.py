"""mc_softmax_ce_fused_static: the softmax_loss kernel with its scratch in a file-scope device array (no
allocation inside the call), through the C ABI, against mc_softmax_ce_fused on the same device inputs.

No counterpart in the reference (added for callers that cannot call cudaMalloc); the arithmetic it must
reproduce is layers.softmax_loss and its VJP (nn/layers.py:201-221; numpy_core.py:170-175).  Collected last on
purpose; passes alone in a fresh process and after the other tests.
"""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_ROWS = 65536        # MC_SOFTMAX_CE_STATIC_MAX_ROWS of include/minpy_cuda.h


def up(a):
    from minpy_b200.cuda_array import CudaArray
    return CudaArray.from_numpy(a)


def empty(shape):
    from minpy_b200.cuda_array import CudaArray
    return CudaArray.empty(shape)


def both(lib, rows, cols, seed):
    rng = np.random.RandomState(seed)
    x = (rng.randn(rows, cols) * 3).astype(np.float32)
    lab = rng.randint(0, cols, rows).astype(np.int32)
    X, L = up(x), up(lab)
    # every buffer exists before the call under test: it may not allocate
    p_ref, dx_ref, loss_ref = empty((rows, cols)), empty((rows, cols)), empty(())
    p_new, dx_new, loss_new = empty((rows, cols)), empty((rows, cols)), empty(())
    lib.mc_softmax_ce_fused(X.ptr, L.ptr, None, p_ref.ptr, dx_ref.ptr, loss_ref.ptr, rows, cols)
    lib.mc_softmax_ce_fused_static(X.ptr, L.ptr, p_new.ptr, dx_new.ptr, loss_new.ptr, rows, cols)
    # the same kernel on the same inputs: bit for bit, the loss summed in the same fixed order
    assert np.array_equal(p_new.asnumpy(), p_ref.asnumpy())
    assert np.array_equal(dx_new.asnumpy(), dx_ref.asnumpy())
    got, want = float(loss_new.asnumpy()), float(loss_ref.asnumpy())
    assert abs(got - want) <= 1e-6 * abs(want), (got, want)
    # float64 restatement (oracle-free: this file must stand alone)
    z = x.astype(np.float64)
    z -= z.max(axis=1, keepdims=True)
    lp = z - np.log(np.exp(z).sum(axis=1, keepdims=True))
    assert abs(got + lp[np.arange(rows), lab].mean()) <= 1e-5 * abs(want)
    # dx only, as the front end calls it
    dx_only, loss_only = empty((rows, cols)), empty(())
    lib.mc_sync()
    free0, free1, total = ctypes.c_size_t(), ctypes.c_size_t(), ctypes.c_size_t()
    lib.mc_mem_info(ctypes.byref(free0), ctypes.byref(total))
    lib.mc_softmax_ce_fused_static(X.ptr, L.ptr, None, dx_only.ptr, loss_only.ptr, rows, cols)
    lib.mc_sync()
    lib.mc_mem_info(ctypes.byref(free1), ctypes.byref(total))
    assert free1.value >= free0.value, 'device memory shrank across the no-allocation call'
    assert np.array_equal(dx_only.asnumpy(), dx_ref.asnumpy())
    assert float(loss_only.asnumpy()) == got


def test_static_scratch_matches_arena_path_two_calls_in_a_row(cuda):
    lib = cuda.L()
    # a large call, then a small one whose row losses sit in front of the large call's leftovers,
    # then the maximum size
    both(lib, 20000, 10, seed=1)
    both(lib, 1024, 10, seed=2)
    both(lib, 7, 3, seed=3)
    both(lib, MAX_ROWS, 32, seed=4)


def test_static_scratch_rejects_sizes_beyond_the_maximum(cuda):
    from minpy_b200._lib import CudaLibError
    lib = cuda.L()
    X, L, out, loss = up(np.zeros((8, 40), np.float32)), up(np.zeros(8, np.int32)), empty((8, 40)), empty(())
    with pytest.raises(CudaLibError, match='32 classes'):
        lib.mc_softmax_ce_fused_static(X.ptr, L.ptr, None, out.ptr, loss.ptr, 8, 40)
    with pytest.raises(CudaLibError, match='MC_SOFTMAX_CE_STATIC_MAX_ROWS'):
        lib.mc_softmax_ce_fused_static(X.ptr, L.ptr, None, out.ptr, loss.ptr, MAX_ROWS + 1, 10)
    with pytest.raises(CudaLibError):
        lib.mc_softmax_ce_fused_static(X.ptr, L.ptr, None, None, loss.ptr, 8, 10)
    # and the array still works afterwards
    both(lib, 256, 10, seed=5)
