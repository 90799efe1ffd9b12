"""Round-2 fused entry points through the C ABI, against (a) the un-fused kernels they replace on the same
device inputs -- bit-exact where the arithmetic is the same -- and (b) the float64 oracle:

  mc_gemm(relu = 2)            signed-zero relu epilogue              nn/layers.py:27,41; numpy_core.py:284-286
  mc_bn_dropout_fwd            batchnorm + dropout mask at the store  nn/layers.py:93-110,141-144
  mc_bn_dropout_relu_bwd       dropout / batchnorm / relu VJPs + bias gradient
  mc_softmax_ce_fused          softmax_loss and its gradient          nn/layers.py:201-221; numpy_core.py:170-175
  mc_avgpool2d_fwd / _bwd      MXNet Pooling(pool_type='avg')
and, through the front end, the lazy forward fusion (`MINPY_FUSE_FORWARD`) of the hidden block of
examples/nn/mlp_bn_dropout.py:41-54: same losses, gradients and running statistics as the un-fused path,
with fewer launches.
"""
import ctypes

import numpy as np
import pytest

from oracle import minpy_oracle as orc

pytestmark = pytest.mark.gpu


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape, (a.shape, b.shape)
    return float(np.max(np.abs(a - b)) / max(1e-30, np.max(np.abs(b)))) if a.size else 0.0


def up(a):
    from minpy_b200.cuda_array import CudaArray
    return CudaArray.from_numpy(a)


def empty(shape):
    from minpy_b200.cuda_array import CudaArray
    return CudaArray.empty(shape)


def launches(lib):
    n = ctypes.c_uint64()
    lib.mc_launch_count(ctypes.byref(n))
    return n.value


@pytest.mark.parametrize('shape', [(64, 96, 32), (256, 512, 128), (1024, 512, 512), (100, 10, 64), (40, 24, 1000)])
@pytest.mark.parametrize('precision', [0, 1])
def test_gemm_signed_zero_relu(cuda, shape, precision):
    """relu mode 2: the values of mode 1 exactly; the sign bit of every output equals (pre-activation < 0)."""
    M, N, K = shape
    lib = cuda.L()
    rng = np.random.RandomState(M + N)
    a = rng.randn(M, K).astype(np.float32)
    b = rng.randn(K, N).astype(np.float32)
    bias = rng.randn(N).astype(np.float32)
    a[3, :] = 0.0                      # a row whose pre-activation is exactly the bias
    bias[1] = 0.0                      # ... and exactly zero in column 1: gradient must pass (x >= 0)
    A, B, bv = up(a), up(b), up(bias)
    outs = []
    for mode in (0, 1, 2):
        C = empty((M, N))
        lib.mc_gemm(0, 0, M, N, K, A.ptr, K, B.ptr, N, C.ptr, N, bv.ptr, mode, 0, precision)
        outs.append(C.asnumpy())
    pre, r1, r2 = outs
    assert np.array_equal(r1, np.maximum(pre, 0))
    assert np.array_equal(r2, r1)                                   # -0.0 == 0.0
    assert np.array_equal(np.signbit(r2), pre < 0)
    assert pre[3, 1] == 0.0 and not np.signbit(r2[3, 1])
    # the VJP factor reads the sign bit
    g = up(rng.randn(M, N).astype(np.float32))
    got = cuda.run_chain(g, 1.0, [(lib.MC_VJP_SIGNPOS_A, up(r2), 0, None, 0, 0.0, 1)], (M, N)).asnumpy()
    assert np.array_equal(got, g.asnumpy() * (pre >= 0))


@pytest.mark.parametrize('N,D', [(1024, 512), (256, 64), (37, 8), (1000, 1000)])
def test_bn_dropout_fwd_equals_the_two_kernels(cuda, N, D):
    lib = cuda.L()
    rng = np.random.RandomState(N)
    x = (rng.randn(N, D) * 2 + 1).astype(np.float32)
    gamma, beta = rng.rand(D).astype(np.float32) + 0.5, rng.randn(D).astype(np.float32)
    rm, rv = rng.randn(D).astype(np.float32), rng.rand(D).astype(np.float32)
    X, G, Bt, RM, RV = up(x), up(gamma), up(beta), up(rm), up(rv)
    p, seed, offset = 0.4, 1234567, 4096
    out1, nrm1, nrv1, sm1, si1 = empty((N, D)), empty((D,)), empty((D,)), empty((D,)), empty((D,))
    lib.mc_batchnorm_train_fwd(X.ptr, G.ptr, Bt.ptr, RM.ptr, RV.ptr, out1.ptr, nrm1.ptr, nrv1.ptr, sm1.ptr, si1.ptr, N, D, 1e-5, 0.9)
    y1 = empty((N, D))
    lib.mc_dropout(out1.ptr, y1.ptr, N * D, p, seed, offset)
    y2, nrm2, nrv2, sm2, si2 = empty((N, D)), empty((D,)), empty((D,)), empty((D,)), empty((D,))
    n0 = launches(lib)
    lib.mc_bn_dropout_fwd(X.ptr, G.ptr, Bt.ptr, RM.ptr, RV.ptr, y2.ptr, nrm2.ptr, nrv2.ptr, sm2.ptr, si2.ptr, N, D, 1e-5, 0.9,
                          p, seed, offset)
    assert launches(lib) - n0 == 1
    for a, b in ((y1, y2), (nrm1, nrm2), (nrv1, nrv2), (sm1, sm2), (si1, si2)):
        assert np.array_equal(a.asnumpy(), b.asnumpy())
    keep = y2.asnumpy() != 0
    if N * D >= 10000:
        assert abs(keep.mean() - (1 - p)) < 0.02
    # oracle: batchnorm (layers.py:93-110) times the device's own mask
    ones = empty((N, D))
    lib.mc_dropout(cuda.full((N, D), 1.0).ptr, ones.ptr, N * D, p, seed, offset)
    mask = ones.asnumpy() > 0.5
    want, _, _ = orc.batchnorm(x.astype(np.float64), gamma.astype(np.float64), beta.astype(np.float64),
                               running_mean=rm.astype(np.float64), running_var=rv.astype(np.float64))
    assert rel(y2.asnumpy(), orc.raw(want) * mask) <= 2e-5


@pytest.mark.parametrize('N,D', [(1024, 512), (256, 64), (37, 8)])
@pytest.mark.parametrize('dropout,relu', [(1, 1), (1, 0), (0, 1), (0, 0)])
def test_bn_dropout_relu_bwd_equals_the_separate_kernels(cuda, N, D, dropout, relu):
    lib = cuda.L()
    rng = np.random.RandomState(N + 2 * dropout + relu)
    pre = rng.randn(N, D).astype(np.float32)
    pre[5, :4] = 0.0                                   # zero pre-activations pass the gradient
    # x = signed-zero relu output of the layer before (what mc_gemm relu = 2 stores)
    x = np.where(pre < 0, np.float32(-0.0), pre).astype(np.float32) if relu else (pre * 2 + 1).astype(np.float32)
    dy = rng.randn(N, D).astype(np.float32)
    gamma = rng.rand(D).astype(np.float32) + 0.5
    X, DY, G = up(x), up(dy), up(gamma)
    out, nrm, nrv, sm, si = empty((N, D)), empty((D,)), empty((D,)), empty((D,)), empty((D,))
    zeros, ones = cuda.zeros((D,)), cuda.ones((D,))
    lib.mc_batchnorm_train_fwd(X.ptr, G.ptr, zeros.ptr, zeros.ptr, ones.ptr, out.ptr, nrm.ptr, nrv.ptr, sm.ptr, si.ptr, N, D, 1e-5, 0.9)
    p, seed, offset = 0.5, 99, 1 << 20
    # the separate kernels
    g1 = DY
    if dropout:
        g1 = empty((N, D))
        lib.mc_dropout(DY.ptr, g1.ptr, N * D, p, seed, offset)
    dx1, dgam1, dbet1 = empty((N, D)), empty((D,)), empty((D,))
    lib.mc_batchnorm_train_bwd(X.ptr, g1.ptr, G.ptr, sm.ptr, si.ptr, dx1.ptr, dgam1.ptr, dbet1.ptr, N, D)
    dx1 = dx1.asnumpy()
    if relu:
        dx1 = dx1 * (pre >= 0)
    # the fused kernel
    dx2, dgam2, dbet2, dbias = empty((N, D)), empty((D,)), empty((D,)), empty((D,))
    n0 = launches(lib)
    lib.mc_bn_dropout_relu_bwd(X.ptr, DY.ptr, G.ptr, sm.ptr, si.ptr, dx2.ptr, dgam2.ptr, dbet2.ptr, dbias.ptr if relu else None,
                               N, D, dropout, p, seed, offset, relu)
    assert launches(lib) - n0 == 1
    assert np.array_equal(dx2.asnumpy(), dx1.astype(np.float32))
    assert np.array_equal(dgam2.asnumpy(), dgam1.asnumpy()) and np.array_equal(dbet2.asnumpy(), dbet1.asnumpy())
    if relu:
        assert rel(dbias.asnumpy(), dx1.astype(np.float64).sum(axis=0)) <= 2e-5
        assert np.all(dx2.asnumpy()[5, :4] == dx1[5, :4])


def test_bn_fused_shape_gate(cuda):
    lib = cuda.L()
    ok = ctypes.c_int()
    for N, D, want in ((1024, 512, 1), (1025, 512, 0), (64, 12, 0), (2048, 512, 0), (1024, 1024, 1), (1024, 1032, 0)):
        lib.mc_bn_fused_supported(N, D, ctypes.byref(ok))
        assert ok.value == want, (N, D)
    X = up(np.zeros((2048, 8), np.float32))
    with pytest.raises(Exception):
        lib.mc_bn_dropout_fwd(X.ptr, X.ptr, X.ptr, X.ptr, X.ptr, X.ptr, X.ptr, X.ptr, X.ptr, X.ptr, 2048, 8, 1e-5, 0.9, 0.5, 1, 0)


@pytest.mark.parametrize('rows,cols', [(1024, 10), (256, 10), (7, 3), (513, 32), (300, 100), (20000, 10)])
def test_softmax_ce_fused_equals_fwd_plus_bwd(cuda, rows, cols):
    lib = cuda.L()
    rng = np.random.RandomState(rows)
    x = (rng.randn(rows, cols) * 3).astype(np.float32)
    lab = rng.randint(0, cols, rows).astype(np.int32)
    X, L = up(x), up(lab)
    probs1, loss1, dx1 = empty((rows, cols)), empty(()), empty((rows, cols))
    lib.mc_softmax_ce_fwd(X.ptr, L.ptr, None, probs1.ptr, loss1.ptr, rows, cols)
    lib.mc_softmax_ce_bwd(probs1.ptr, L.ptr, None, dx1.ptr, rows, cols)
    probs2, loss2, dx2 = empty((rows, cols)), empty(()), empty((rows, cols))
    n0 = launches(lib)
    lib.mc_softmax_ce_fused(X.ptr, L.ptr, None, probs2.ptr, dx2.ptr, loss2.ptr, rows, cols)
    if cols <= 32:
        assert launches(lib) - n0 == 1
    assert rel(probs2.asnumpy(), probs1.asnumpy()) <= 5e-6 and rel(dx2.asnumpy(), dx1.asnumpy()) <= 5e-6
    assert abs(float(loss2.asnumpy()) - float(loss1.asnumpy())) <= 1e-6 * abs(float(loss1.asnumpy()))
    # oracle
    want = orc.softmax_loss(x.astype(np.float64), lab)
    assert abs(float(loss2.asnumpy()) - float(orc.raw(want))) <= 1e-5 * abs(float(orc.raw(want)))
    e = np.exp(x.astype(np.float64) - x.max(axis=1, keepdims=True))
    pr = e / e.sum(axis=1, keepdims=True)
    pr[np.arange(rows), lab] -= 1
    assert rel(dx2.asnumpy(), pr / rows) <= 1e-5
    # dx only (what the front end asks for) and one-hot labels
    dx3, loss3 = empty((rows, cols)), empty(())
    lib.mc_softmax_ce_fused(X.ptr, L.ptr, None, None, dx3.ptr, loss3.ptr, rows, cols)
    assert np.array_equal(dx3.asnumpy(), dx2.asnumpy())
    hot = np.zeros((rows, cols), np.float32)
    hot[np.arange(rows), lab] = 1
    dx4, loss4 = empty((rows, cols)), empty(())
    lib.mc_softmax_ce_fused(X.ptr, None, up(hot).ptr, None, dx4.ptr, loss4.ptr, rows, cols)
    assert rel(dx4.asnumpy(), dx1.asnumpy()) <= 5e-6


@pytest.mark.parametrize('shape,k,stride', [((4, 6, 8, 8), 8, 1), ((2, 3, 9, 7), 2, 2), ((3, 5, 10, 10), 3, 1), ((2, 4, 7, 7), 3, 2)])
def test_avgpool_matches_torch_float64(cuda, shape, k, stride):
    import torch
    rng = np.random.RandomState(k)
    x = rng.randn(*shape).astype(np.float32)
    y = cuda.avgpool2d_fwd(up(x), k, stride)
    t = torch.tensor(x.astype(np.float64), requires_grad=True)
    ty = torch.nn.functional.avg_pool2d(t, k, stride)
    assert rel(y.asnumpy(), ty.detach().numpy()) <= 1e-6
    g = rng.randn(*y.shape).astype(np.float32)
    ty.backward(torch.tensor(g.astype(np.float64)))
    dx = cuda.avgpool2d_bwd(up(g), shape, k, stride)
    assert rel(dx.asnumpy(), t.grad.numpy()) <= 1e-6


def _block_step(fuse, graph, drop=0.5):
    """Three steps of a 2-hidden-layer affine-relu-batchnorm-dropout net (mlp_bn_dropout.py:41-54 block)."""
    import importlib
    import minpy
    minpy.set_global_policy('only_cuda')
    from minpy.array_variants.cuda import cuda_core
    import minpy.numpy.random as mrandom
    from minpy import array as marray
    from minpy.nn.io import DataBatch
    from minpy.nn.solver import Solver
    from minpy_b200 import workloads
    from minpy_b200.cuda_array import CudaArray
    from minpy_b200._lib import get_lib
    saved = cuda_core._FUSE_FORWARD  # pylint: disable=protected-access
    cuda_core._FUSE_FORWARD = fuse  # pylint: disable=protected-access
    try:
        mrandom.seed(21)
        rng = np.random.RandomState(4)
        model = workloads.DeepMLP(in_dim=96, hidden=(64, 64), classes=10, drop=drop)
        solver = Solver(model, None, None, update_rule='sgd_momentum', optim_config={'learning_rate': 0.05, 'momentum': 0.9},
                        init_rule='gaussian', init_config={'stdvar': 0.1}, verbose=False, step_graph=graph)
        solver.init()
        for k, v in workloads.init_params(model, 3, scale=0.1).items():
            model.params[k] = marray.wrap(CudaArray.from_numpy(v))
        lib = get_lib()
        counts = []
        for i in range(4):
            X = rng.randn(128, 96).astype(np.float32)
            Y = rng.randint(0, 10, 128)
            n0 = launches(lib)
            solver._step(DataBatch([marray.wrap(CudaArray.from_numpy(X))], [marray.wrap(CudaArray.from_numpy(Y))], pad=0, index=i))  # pylint: disable=protected-access
            counts.append(launches(lib) - n0)
        params = dict((k, v.asnumpy()) for k, v in model.params.items())
        aux = dict((k, v.asnumpy()) for k, v in model.aux_params.items())
        losses = [float(l) for l in solver.loss_history]
        return params, aux, losses, counts
    finally:
        cuda_core._FUSE_FORWARD = saved  # pylint: disable=protected-access
        minpy.set_global_policy('prefer_cuda')


def test_forward_fusion_equals_the_unfused_step_with_fewer_launches():
    fused = _block_step(True, False)
    plain = _block_step(False, False)
    for a, b in zip(fused[:2], plain[:2]):
        for k in a:
            assert rel(a[k], b[k]) <= 1e-4, k          # (bias gradients are summed in another order: 2e-5 after 4 steps)
    assert np.allclose(fused[2], plain[2], rtol=1e-5)
    # eager launches per step: 2 hidden blocks save >= 2 forward and >= 3 backward launches each, the loss 3
    assert fused[3][-1] <= plain[3][-1] - 8, (fused[3], plain[3])


def test_forward_fusion_inside_the_step_graph():
    """The lazily evaluated values live and die inside the capture; replays equal eager stepping.  (p = 0:
    a replay draws its masks from a shifted Philox offset, so only a keep-everything dropout is comparable.)"""
    eager = _block_step(True, False, drop=0.0)
    graph = _block_step(True, True, drop=0.0)
    for a, b in zip(eager[:2], graph[:2]):
        for k in a:
            assert rel(a[k], b[k]) <= 1e-5, k
    assert np.allclose(eager[2], graph[2], rtol=1e-5)


def test_lazy_values_evaluate_on_demand():
    """A pending affine / batchnorm result that is NOT followed by relu / dropout evaluates when touched."""
    import minpy
    minpy.set_global_policy('only_cuda')
    try:
        import minpy.numpy as mnp
        from minpy.nn import layers
        from minpy.array_variants import ArrayType
        from minpy_b200.lazy import LazyArray
        rng = np.random.RandomState(0)
        x, w, b = rng.randn(32, 16), rng.randn(16, 8), rng.randn(8)
        a = layers.affine(mnp.array(x), mnp.array(w), mnp.array(b))
        raw = a.get_data(ArrayType.CUDA)
        assert isinstance(raw, LazyArray) and raw.pending() and a.shape == (32, 8)
        h = layers.relu(a)
        assert raw.pending()                                  # the relu ran the GEMM itself
        assert rel(h.asnumpy(), np.maximum(x.dot(w) + b, 0)) <= 1e-2
        assert rel(a.asnumpy(), x.dot(w) + b) <= 1e-2 and not raw.pending()      # ... and the pre-activation on demand
        out, rm, rv = layers.batchnorm(h, mnp.ones((8,)), mnp.zeros((8,)))
        assert out.get_data(ArrayType.CUDA).pending()
        y = layers.dropout(out, 0.5)
        assert out.get_data(ArrayType.CUDA).pending() and not rm.get_data(ArrayType.CUDA).pending()
        hn = np.maximum(x.dot(w) + b, 0)
        bn = (hn - hn.mean(0)) / np.sqrt(hn.var(0) + 1e-5)
        yv = y.asnumpy()
        assert rel(yv[yv != 0], bn[yv != 0]) <= 2e-2
        assert rel(out.asnumpy(), bn) <= 2e-2                 # the un-dropped batchnorm output, on demand
        assert rel(rm.asnumpy(), 0.1 * hn.mean(0)) <= 2e-2
    finally:
        minpy.set_global_policy('prefer_cuda')


@pytest.mark.parametrize('N,T,V', [(128, 100, 128), (4, 7, 10), (16, 5, 300), (3, 2, 33)])
def test_temporal_softmax_loss_matches_the_oracle(N, T, V):
    """mc_temporal_softmax_loss through layers.temporal_softmax_loss: value and gradient equal the oracle's
    composition (nn/layers.py:444-480 differentiated by its tape), including masked-out steps and a
    non-unit upstream gradient."""
    import minpy
    minpy.set_global_policy('only_cuda')
    try:
        import minpy.numpy as mnp
        from minpy import core
        from minpy.nn import layers
        rng = np.random.RandomState(N + T)
        x = (rng.randn(N, T, V) * 2).astype(np.float32)
        y = rng.randint(0, V, (N, T))
        mask = (rng.rand(N, T) > 0.3).astype(np.float32)
        yd, md = mnp.array(y), mnp.array(mask)
        g, loss = core.grad_and_loss(lambda v: layers.temporal_softmax_loss(v, yd, md) * 3.0)(mnp.array(x))
        og, oloss = orc.grad_and_loss(lambda v: orc.temporal_softmax_loss(v, y, mask.astype(np.float64)) * 3.0)(x.astype(np.float64))
        assert abs(float(loss) - float(oloss)) <= 2e-5 * abs(float(oloss))
        assert rel(g.asnumpy(), orc.raw(og)) <= 2e-5
        assert np.all(g.asnumpy()[mask == 0] == 0)
    finally:
        minpy.set_global_policy('prefer_cuda')


def test_lstm_fused_step_opt_in_matches_default():
    """MC_LSTM_FUSED=1 (recurrent GEMM with the gate arithmetic as its cluster epilogue) gives the same
    lstm_temporal value and gradients as the default GEMM + gate-kernel path."""
    import os, subprocess, sys
    code = r"""
import numpy as np, minpy
minpy.set_global_policy('only_cuda')
import minpy.numpy as mnp
from minpy import core
from minpy.nn import layers
rng = np.random.RandomState(0)
N, T, D, H = 128, 6, 64, 256
x = rng.randn(N, T, D).astype(np.float32); h0 = rng.randn(N, H).astype(np.float32) * 0.1
Wx = (rng.randn(D, 4 * H) * 0.1).astype(np.float32); Wh = (rng.randn(H, 4 * H) * 0.1).astype(np.float32); b = rng.randn(4 * H).astype(np.float32) * 0.1
f = lambda x_, h_, wx, wh, b_: mnp.sum(layers.lstm_temporal(x_, h_, wx, wh, b_) ** 2)
g, v = core.grad_and_loss(f, argnum=range(5))(*[mnp.array(a) for a in (x, h0, Wx, Wh, b)])
np.savez(__import__('sys').argv[1], v=float(v), **dict(('g%d' % i, a.asnumpy()) for i, a in enumerate(g)))
"""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    outs = []
    for flag in ('0', '1'):
        path = os.path.join(root, 'gpurun_out', 'lstm_fused_%s.npz' % flag)
        os.makedirs(os.path.dirname(path), exist_ok=True)
        res = subprocess.run([sys.executable, '-c', code, path], env=dict(os.environ, PYTHONPATH=root, MC_LSTM_FUSED=flag),
                             cwd=root, capture_output=True, text=True, timeout=300)
        assert res.returncode == 0, res.stderr[-3000:]
        outs.append(dict(np.load(path)))
    a, b = outs
    assert abs(a['v'] - b['v']) <= 2e-3 * abs(a['v'])
    for k in a:
        if k != 'v':
            assert rel(b[k], a[k]) <= 5e-3, k


@pytest.mark.parametrize('rows,K,C', [(1024, 512, 10), (256, 512, 10), (512, 2048, 10), (37, 50, 3), (128, 96, 16),
                                      (9, 2600, 7), (64, 3072, 16), (5, 1, 1), (100, 513, 5), (33, 1000, 12)])
def test_classifier_head_kernel_matches_oracle_and_composition(cuda, rows, K, C):
    """mc_head_softmax_loss: logits, loss, dlogits and dh against float64 (affine + softmax_loss differentiated by
    the oracle's tape) and, through the front end, against the separate kernels (MINPY_FUSE_FORWARD off)."""
    lib = cuda.L()
    rng = np.random.RandomState(rows + K)
    h = rng.randn(rows, K).astype(np.float32)
    W = (rng.randn(K, C) * 0.1).astype(np.float32)
    b = rng.randn(C).astype(np.float32)
    lab = rng.randint(0, C, rows).astype(np.int32)
    H, Wd, B, L = up(h), up(W), up(b), up(lab)
    logits, dl, dh, loss = empty((rows, C)), empty((rows, C)), empty((rows, K)), empty(())
    n0 = launches(lib)
    lib.mc_head_softmax_loss(H.ptr, Wd.ptr, B.ptr, L.ptr, logits.ptr, dl.ptr, dh.ptr, loss.ptr, rows, K, C)
    assert launches(lib) - n0 == 1
    f = lambda hv, wv, bv: orc.softmax_loss(orc.affine(hv, wv, bv), lab)
    (gh, gw, gb), want = orc.grad_and_loss(f, argnum=range(3))(h.astype(np.float64), W.astype(np.float64), b.astype(np.float64))
    assert rel(logits.asnumpy(), h.astype(np.float64).dot(W) + b) <= 1e-5
    assert abs(float(loss.asnumpy()) - float(orc.raw(want))) <= 1e-5 * abs(float(orc.raw(want)))
    assert rel(dh.asnumpy(), orc.raw(gh)) <= 2e-5
    assert rel(dl.asnumpy().sum(axis=0), orc.raw(gb)) <= 2e-5
    assert rel(h.T.astype(np.float64).dot(dl.asnumpy()), orc.raw(gw)) <= 2e-5
    # front end: the same head through layers.affine + layers.softmax_loss, fused and un-fused
    import minpy
    minpy.set_global_policy('only_cuda')
    try:
        import minpy.numpy as mnp
        from minpy import core
        from minpy.nn import layers
        from minpy.array_variants.cuda import cuda_core
        res = []
        for fuse in (True, False):
            saved, cuda_core._FUSE_FORWARD = cuda_core._FUSE_FORWARD, fuse          # pylint: disable=protected-access
            try:
                g, v = core.grad_and_loss(lambda hv, wv, bv: layers.softmax_loss(layers.affine(hv, wv, bv), mnp.array(lab)),
                                          argnum=range(3))(mnp.array(h), mnp.array(W), mnp.array(b))
                res.append((float(v), [a.asnumpy() for a in g]))
            finally:
                cuda_core._FUSE_FORWARD = saved                                     # pylint: disable=protected-access
        assert abs(res[0][0] - res[1][0]) <= 1e-5 * abs(res[1][0])
        # dW = h^T . dlogits is a tcgen05 TF32 GEMM on the chain when the shape is large enough (mc_gemm routing)
        for a, b2, o, tol in zip(res[0][1], res[1][1], (gh, gw, gb), (2e-5, 2e-3, 2e-5)):
            assert rel(a, orc.raw(o)) <= tol and rel(b2, orc.raw(o)) <= tol
    finally:
        minpy.set_global_policy('prefer_cuda')


@pytest.mark.parametrize('N,C,H,F,K,pad', [(8, 3, 32, 32, 3, 1), (4, 32, 16, 64, 3, 1), (4, 64, 8, 128, 3, 1), (6, 3, 32, 32, 7, 0),
                                           (3, 16, 12, 32, 3, 1), (2, 8, 20, 64, 5, 2), (5, 4, 6, 32, 3, 1)])
def test_conv_relu_pool_epilogue_equals_conv_then_pool(cuda, N, C, H, F, K, pad):
    """mc_conv2d_nhwc_fwd_pool against mc_conv2d_nhwc_fwd (+relu) followed by mc_maxpool2d_nhwc_fwd: the same MMAs,
    so pooled values AND window indices must be bit-identical; and the symbol path with / without the fusion gives
    the same loss-side gradients (the pooling VJP consumes the index either way)."""
    rng = np.random.RandomState(N * 100 + H)
    x = rng.randn(N, C, H, H).astype(np.float32)
    w = (rng.randn(F, C, K, K) * 0.2).astype(np.float32)
    b = rng.randn(F).astype(np.float32) * 0.1
    x[0, :, :2, :2] = 0.0                       # a window of exact ties
    a = cuda.to_nhwc(up(x), pad if cuda.conv_merged_ok(C, K) else 0)
    W, B = up(w), up(b)
    assert cuda.conv_pool_fusable(a, W, pad)
    full = cuda.conv2d_nhwc_fwd(a, W, B, pad, relu=True)
    y1, i1 = cuda.maxpool2d_nhwc_fwd(full, 2, want_idx=True)
    y2, i2, like = cuda.conv2d_nhwc_fwd_pool(a, W, B, pad, relu=True, want_idx=True)
    assert (like.N, like.C, like.H, like.W) == (full.N, full.C, full.H, full.W)
    assert np.array_equal(y1.data.asnumpy(), y2.data.asnumpy())
    assert np.array_equal(i1.asnumpy(), i2.asnumpy())
    y3, i3, _ = cuda.conv2d_nhwc_fwd_pool(a, W, B, pad, relu=True, want_idx=False)
    assert i3 is None and np.array_equal(y3.data.asnumpy(), y1.data.asnumpy())
    # against the float64 oracle (TF32 contraction: 1e-2 of the largest entry)
    want = orc.raw(orc.maxpool2d(orc.relu_mx(orc.conv2d(x.astype(np.float64), w.astype(np.float64), b.astype(np.float64), 1, pad)), 2, 2))
    assert rel(cuda.to_nchw(y2).asnumpy(), want) <= 1e-2


def test_conv_pool_fusion_in_the_symbol_path_matches_unfused():
    """ConvNet C4 in miniature through core.Function: gradients with MINPY_CONV_POOL_FUSE on and off are equal."""
    import minpy
    minpy.set_global_policy('only_cuda')
    try:
        import minpy.numpy as mnp
        from minpy import core
        from minpy_b200 import workloads, symbol_exec
        rng = np.random.RandomState(5)
        X = rng.randn(16, 3, 32, 32).astype(np.float32)
        Y = rng.randint(0, 10, 16)
        res = []
        for fuse in (True, False):
            saved, symbol_exec._POOL_FUSE = symbol_exec._POOL_FUSE, fuse       # pylint: disable=protected-access
            try:
                model = workloads.ConvNet(batch=16)
                init = workloads.init_params(model, 3, scale=0.1)
                names = sorted(init)
                for k in names:
                    model.params[k] = mnp.array(init[k])

                def loss(*ps):
                    for k, pv in zip(names, ps):
                        model.params[k] = pv
                    return model.loss(model.forward(mnp.array(X), 'train'), mnp.array(Y))
                g, v = core.grad_and_loss(loss, argnum=range(len(names)))(*[model.params[k] for k in names])
                res.append((float(v), [a.asnumpy() for a in g]))
            finally:
                symbol_exec._POOL_FUSE = saved                                  # pylint: disable=protected-access
        assert res[0][0] == res[1][0]
        for a, b in zip(res[0][1], res[1][1]):
            assert np.array_equal(a, b)
    finally:
        minpy.set_global_policy('prefer_cuda')
