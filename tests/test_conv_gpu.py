"""GPU parity of the convolution kernels (C ABI: mc_conv2d_fwd / dgrad / wgrad, mc_maxpool2d_*)
against the oracle's NumPy restatement (oracle/minpy_oracle.py conv2d_* -- "parity unpinned":
MXNet is an un-vendored dependency of the reference).

Two paths behind one entry point are both held to the oracle:
  * precision 1: implicit GEMM on tcgen05 (kind::tf32, TMA-gathered windows) -- 1e-2 relative
    (north_star's TF32 tolerance), and the entry point must report that it took that path;
  * precision 0: direct FP32 CUDA-core kernels (exact accumulate) -- 2e-5 relative; this is the
    FP32-accumulate cross-check of the tensor-core path.
"""
import ctypes

import numpy as np
import pytest

from oracle import minpy_oracle as orc

pytestmark = pytest.mark.gpu


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape, (a.shape, b.shape)
    return float(np.max(np.abs(a - b)) / max(1e-30, np.max(np.abs(b))))


# (N, C, H, W, F, K, pad)
SHAPES = [
    (4, 3, 32, 32, 32, 3, 1),     # C4 layer 1: C = 3 (one 8-channel k-block, 5 zero-filled)
    (4, 32, 16, 16, 64, 3, 1),    # C4 layer 2: 16-wide rows, two rows per 32-pixel chunk
    (4, 64, 8, 8, 128, 3, 1),     # C4 layer 3: 8-wide rows, four rows per chunk
    (3, 3, 32, 32, 32, 7, 0),     # examples/nn/cnn.py: 7x7, 26-wide output on a 32-wide virtual row
    (2, 16, 32, 32, 48, 5, 2),    # F not a power of two
    (2, 8, 64, 64, 16, 3, 1),     # two chunks per row
    (5, 40, 16, 16, 24, 3, 1),    # C not a multiple of the 32-channel k-block
    (2, 4, 12, 12, 8, 3, 1),      # 12-wide rows on a 16-wide virtual row
    (1, 3, 8, 8, 4, 3, 1),        # a single partial tile
    (2, 3, 20, 20, 8, 3, 0),      # 18-wide output
    (3, 130, 8, 8, 260, 1, 0),    # 1x1, more than one channel block / output tile
    (2, 3, 9, 9, 5, 3, 1),        # odd width, F not a multiple of 4 (pitch padding)
    (2, 6, 10, 14, 7, 3, 2),      # pad 2 with a 3x3 filter: output larger than input
]


def last_path(lib):
    p = ctypes.c_int(-1)
    lib.mc_conv_last_path(ctypes.byref(p))
    return p.value


# TF32 tolerance ~3x the measured worst case (gpurun_out/parity_errors.json), inside north_star's 1e-2
@pytest.mark.parametrize('mode,tol', [('tf32', 4e-3), ('fp32', 2e-5)])
@pytest.mark.parametrize('shape', SHAPES)
def test_conv_fwd_dgrad_wgrad(cuda, shape, mode, tol):
    N, C, H, W, F, K, pad = shape
    lib = cuda.L()
    rng = np.random.RandomState(sum(shape))
    x = rng.randn(N, C, H, W).astype(np.float32)
    w = (rng.randn(F, C, K, K) / np.sqrt(C * K * K)).astype(np.float32)
    b = rng.randn(F).astype(np.float32)
    from minpy_b200.cuda_array import CudaArray
    xd, wd, bd = CudaArray.from_numpy(x), CudaArray.from_numpy(w), CudaArray.from_numpy(b)
    x64, w64, b64 = x.astype(np.float64), w.astype(np.float64), b.astype(np.float64)
    cuda.set_gemm_precision(mode)
    try:
        y = cuda.conv2d_fwd(xd, wd, bd, 1, pad, relu=False)
        path_f = last_path(lib)
        yr = cuda.conv2d_fwd(xd, wd, bd, 1, pad, relu=True)
        ref = orc.conv2d_fwd(x64, w64, b64, 1, pad)
        from conftest import record_err
        assert record_err('conv_fwd_' + mode, rel(y.asnumpy(), ref)) <= tol
        assert rel(yr.asnumpy(), np.maximum(ref, 0)) <= tol
        dy = rng.randn(*ref.shape).astype(np.float32)
        dx, dw, db = cuda.conv2d_bwd(xd, wd, CudaArray.from_numpy(dy), 1, pad)
        path_w = last_path(lib)
        rdx, rdw, rdb = orc.conv2d_bwd(x64, w64, dy.astype(np.float64), 1, pad)
        assert record_err('conv_dgrad_' + mode, rel(dx.asnumpy(), rdx)) <= tol
        assert record_err('conv_wgrad_' + mode, rel(dw.asnumpy(), rdw)) <= tol
        # tf32 mode: the bias gradient is one more accumulator of the wgrad kernel (dy x ones on the
        # tensor cores), so dy is rounded to TF32 there as well
        assert rel(db.asnumpy(), rdb) <= (tol if mode == 'tf32' else 2e-5)
        _, dw2, _ = cuda.conv2d_bwd(xd, wd, CudaArray.from_numpy(dy), 1, pad, need_dx=False, need_db=False)
        assert np.array_equal(dw2.asnumpy(), dw.asnumpy())       # deterministic split reduction
    finally:
        cuda.set_gemm_precision('tf32')
    if mode == 'tf32':
        assert path_f == 1 and path_w == 1, 'the tensor-core path must serve every stride-1 convolution'
    else:
        assert path_f == 0 and path_w == 0


def test_conv_strided_falls_back_to_direct_kernels(cuda):
    """stride 2 is served by the direct FP32 kernels (the tensor-core path is stride 1)."""
    lib = cuda.L()
    from minpy_b200.cuda_array import CudaArray
    rng = np.random.RandomState(5)
    for (N, C, H, W, F, K, pad, stride) in [(2, 4, 16, 16, 6, 3, 1, 2), (2, 3, 9, 9, 4, 3, 0, 2)]:
        x = rng.randn(N, C, H, W).astype(np.float32)
        w = rng.randn(F, C, K, K).astype(np.float32) * 0.2
        b = rng.randn(F).astype(np.float32)
        y = cuda.conv2d_fwd(CudaArray.from_numpy(x), CudaArray.from_numpy(w), CudaArray.from_numpy(b), stride, pad)
        assert last_path(lib) == 0
        ref = orc.conv2d_fwd(x.astype(np.float64), w.astype(np.float64), b.astype(np.float64), stride, pad)
        assert rel(y.asnumpy(), ref) <= 2e-5
        dy = rng.randn(*ref.shape).astype(np.float32)
        dx, dw, db = cuda.conv2d_bwd(CudaArray.from_numpy(x), CudaArray.from_numpy(w), CudaArray.from_numpy(dy), stride, pad)
        rdx, rdw, rdb = orc.conv2d_bwd(x.astype(np.float64), w.astype(np.float64), dy.astype(np.float64), stride, pad)
        assert rel(dx.asnumpy(), rdx) <= 2e-5 and rel(dw.asnumpy(), rdw) <= 2e-5 and rel(db.asnumpy(), rdb) <= 2e-5


def test_conv_linearity_at_full_c4_size(cuda):
    """Size-independent property at the BASELINE config size (batch 512, 32x32, 3 -> 32):
    conv(a*x1 + x2) == a*conv(x1) + conv(x2) within TF32 rounding; first / last image agree with
    the oracle evaluated on just those images."""
    from minpy_b200.cuda_array import CudaArray
    rng = np.random.RandomState(1)
    N = 512
    x1 = rng.randn(N, 3, 32, 32).astype(np.float32)
    x2 = rng.randn(N, 3, 32, 32).astype(np.float32)
    w = (rng.randn(32, 3, 3, 3) * 0.2).astype(np.float32)
    wd = CudaArray.from_numpy(w)
    y1 = cuda.conv2d_fwd(CudaArray.from_numpy(x1), wd, None, 1, 1).asnumpy()
    y2 = cuda.conv2d_fwd(CudaArray.from_numpy(x2), wd, None, 1, 1).asnumpy()
    y12 = cuda.conv2d_fwd(CudaArray.from_numpy(2.0 * x1 + x2), wd, None, 1, 1).asnumpy()
    assert rel(y12, 2.0 * y1 + y2) <= 1e-2
    for i in (0, N - 1):
        ref = orc.conv2d_fwd(x1[i:i + 1].astype(np.float64), w.astype(np.float64), None, 1, 1)
        assert rel(y1[i:i + 1], ref) <= 1e-2


@pytest.mark.parametrize('shape', [(6, 5, 8, 8, 2, 2), (3, 4, 13, 13, 2, 2), (2, 3, 9, 9, 3, 2), (2, 2, 6, 6, 2, 1)])
def test_maxpool_fwd_bwd(cuda, shape):
    N, C, H, W, k, s = shape
    from minpy_b200.cuda_array import CudaArray
    rng = np.random.RandomState(3)
    x = rng.randn(N, C, H, W).astype(np.float32)
    y, arg = cuda.maxpool2d_fwd(CudaArray.from_numpy(x), k, s)
    ref, rarg = orc.maxpool2d_fwd(x.astype(np.float64), k, s)
    assert np.array_equal(y.asnumpy(), ref.astype(np.float32))          # bit-exact: pooling only moves values
    assert np.array_equal(arg.asnumpy().astype(np.int64), rarg)         # arg-max index bit-exact
    dy = rng.randn(*ref.shape).astype(np.float32)
    dx = cuda.maxpool2d_bwd(CudaArray.from_numpy(dy), arg, x.shape)
    assert rel(dx.asnumpy(), orc.maxpool2d_bwd(dy.astype(np.float64), rarg, x.shape)) <= 1e-6


# ------------------------------------------------------------- channels-last building blocks --
@pytest.mark.parametrize('shape,wpad', [((3, 3, 8, 8), 1), ((2, 5, 7, 9), 0), ((2, 64, 6, 6), 0), ((4, 3, 32, 32), 3),
                                        ((1, 33, 5, 40), 2)])
def test_nhwc_layout_round_trip_is_bit_exact(cuda, shape, wpad):
    from minpy_b200.cuda_array import CudaArray
    rng = np.random.RandomState(2)
    x = rng.randn(*shape).astype(np.float32)
    a = cuda.to_nhwc(CudaArray.from_numpy(x), wpad)
    N, C, H, W = shape
    h = a.data.asnumpy()
    assert h.shape == (N, H, W + 2 * wpad, (C + 3) // 4 * 4)
    assert np.array_equal(h[:, :, wpad:wpad + W, :C], x.transpose(0, 2, 3, 1))
    assert not h[:, :, :wpad].any() and not h[:, :, wpad + W:].any() and not h[..., C:].any()    # pads are zero
    if not wpad:
        assert np.array_equal(cuda.to_nchw(a).asnumpy(), x)


@pytest.mark.parametrize('shape,k', [((4, 32, 16, 16), 2), ((2, 6, 12, 12), 3), ((3, 8, 9, 7), 2), ((2, 3, 8, 8), 2)])
def test_maxpool_nhwc_matches_oracle_and_folds_relu_mask(cuda, shape, k):
    from minpy_b200.cuda_array import CudaArray
    rng = np.random.RandomState(4)
    x = np.maximum(rng.randn(*shape), 0).astype(np.float32)      # a relu'd tensor: exact zeros, ties in windows
    a = cuda.to_nhwc(CudaArray.from_numpy(x))
    y, idx = cuda.maxpool2d_nhwc_fwd(a, k)
    ref, rarg = orc.maxpool2d_fwd(x.astype(np.float64), k, k)
    assert np.array_equal(cuda.to_nchw(y).asnumpy(), ref.astype(np.float32))
    dy = rng.randn(*ref.shape).astype(np.float32)
    dyh = cuda.to_nhwc(CudaArray.from_numpy(dy))
    dx = cuda.to_nchw(cuda.maxpool2d_nhwc_bwd(dyh, idx, a, k)).asnumpy()
    assert np.array_equal(dx, orc.maxpool2d_bwd(dy.astype(np.float64), rarg, x.shape).astype(np.float32))   # same arg-max
    dxm = cuda.to_nchw(cuda.maxpool2d_nhwc_bwd(dyh, idx, a, k, relu_y=y)).asnumpy()
    want = orc.maxpool2d_bwd((dy * (ref > 0)).astype(np.float64), rarg, x.shape).astype(np.float32)
    assert np.array_equal(dxm, want)                               # == pool bwd followed by the relu mask out > 0
