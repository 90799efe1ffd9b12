"""Host logic of the lazy forward fusion and the VJP peepholes (minpy_b200/lazy.py, cuda_core.py fused_*),
checked WITHOUT a GPU: the library binding is replaced by a recorder and device buffers by fakes, and the
test asserts WHICH library calls the hidden block affine -> relu -> batchnorm -> dropout of
examples/nn/mlp_bn_dropout.py:41-54 issues, forward and backward.  Numerics are the GPU suite's job
(tests/test_fused_gpu.py); this pins the call sequence and the fallbacks."""
import numpy as np
import pytest


class _Lib(object):
    """Records every mc_* call; enums come from the real header."""

    def __init__(self, enums):
        self.calls = []
        self.enums = enums
        for k, v in enums.items():
            setattr(self, k, v)

    def __getattr__(self, name):
        if not name.startswith('mc_'):
            raise AttributeError(name)

        def call(*args):
            self.calls.append((name, args))
            return 0
        call.raw = call
        return call

    def names(self):
        return [c[0] for c in self.calls]


@pytest.fixture
def fake(monkeypatch):
    from minpy_b200 import ops, cuda_array, _lib
    from minpy_b200.cuda_array import _Buffer, _state
    lib = _Lib(_lib.parse_enums())
    counter = [4096]

    def fake_init(self, nbytes):
        self.nbytes = (nbytes + 511) & ~511 if nbytes else 512
        self.device, self.pool = 0, None
        self.ptr = counter[0]
        counter[0] += self.nbytes

    monkeypatch.setattr(_Buffer, '__init__', fake_init)
    monkeypatch.setattr(_Buffer, '__del__', lambda self: None)
    monkeypatch.setattr(ops, 'L', lambda: lib)
    monkeypatch.setattr(ops, '_lib', lib, raising=False)
    monkeypatch.setitem(_state, 'device', 0)
    monkeypatch.setitem(_state, 'home', 0)
    monkeypatch.setattr(cuda_array, 'get_lib', lambda: lib)
    ops._side.update(keep=[], pending=False, active=False)      # pylint: disable=protected-access
    yield lib
    ops._side.update(keep=[], pending=False, active=False)      # pylint: disable=protected-access


def _arr(shape):
    from minpy_b200.cuda_array import CudaArray
    return CudaArray.empty(shape)


def test_hidden_block_forward_and_backward_call_sequence(fake):
    from minpy.array_variants.cuda import cuda_core as cc
    from minpy_b200.lazy import LazyArray, ChainGrad, DROPOUT
    N, D, H = 128, 96, 64
    x, w, b = _arr((N, D)), _arr((D, H)), _arr((H,))
    gamma, beta, rm, rv = _arr((H,)), _arr((H,)), _arr((H,)), _arr((H,))

    a = cc.fused_affine(x, w, b)
    assert isinstance(a, LazyArray) and a.pending() and fake.calls == []
    h = cc.fused_relu(a)
    assert fake.names() == ['mc_gemm'] and fake.calls[0][1][12] == 2          # relu mode 2 in the epilogue
    assert a.pending() and h._tag is cc._RELU_SIGNED                          # pylint: disable=protected-access
    out, nrm, nrv = cc.fused_batchnorm(h, gamma, beta, rm, rv, 1e-5, 0.9)
    assert out.pending() and nrm.pending() and fake.names() == ['mc_gemm']
    y = cc.fused_dropout(out, 0.5, 7, 1024)
    assert fake.names() == ['mc_gemm', 'mc_bn_dropout_fwd']
    assert out.pending() and not nrm.pending() and not nrv.pending()
    assert fake.calls[1][1][-3:] == (0.5, 7, 1024)

    # ---- backward, in tape order: dropout, batchnorm, relu, affine ----
    del fake.calls[:]
    g_y = _arr((N, H))
    g_out = cc._dropout_vjp(y, out, 0.5, 7, 1024)(g_y)                        # pylint: disable=protected-access
    assert isinstance(g_out, ChainGrad) and g_out.steps[0][0] == DROPOUT and fake.calls == []
    dh, dgamma, dbeta = cc._batchnorm_vjp((out, nrm, nrv), h, gamma, beta, rm, rv, 1e-5, 0.9)((g_out, None, None))  # pylint: disable=protected-access
    assert fake.names() == ['mc_bn_dropout_relu_bwd']
    args = fake.calls[0][1]
    assert args[0] == h.ptr and args[1] == g_y.ptr and args[8] is not None     # x, dy, dbias
    assert args[11:] == (1, 0.5, 7, 1024, 1)                                   # dropout on, (p, seed, offset), relu on
    assert dh._tag[0] == 'relu_done' and dh._tag[1] is h                       # pylint: disable=protected-access
    g_a = cc._relu_vjp(h, a)(dh)                                               # pylint: disable=protected-access
    assert g_a is dh and len(fake.calls) == 1                                  # relu already applied: pass-through
    dx_fn, dw_fn, db_fn = cc._affine_vjps()                                    # pylint: disable=protected-access
    db = db_fn(a, x, w, b)(g_a)
    assert db.ptr == dh._tag[2].ptr and len(fake.calls) == 1                   # pylint: disable=protected-access
    dw_fn(a, x, w, b)(g_a)
    dx_fn(a, x, w, b)(g_a)
    # the weight gradient runs on the side stream, the data gradient stays on the compute stream
    assert fake.names() == ['mc_bn_dropout_relu_bwd', 'mc_side_begin', 'mc_gemm', 'mc_side_end', 'mc_gemm']
    from minpy_b200 import ops
    assert ops._side['pending'] and any(k is x for k in ops._side['keep'])     # pylint: disable=protected-access
    ops.side_join()
    assert fake.names()[-1] == 'mc_side_join' and not ops._side['keep']        # pylint: disable=protected-access
    assert a.pending() and out.pending()                                       # never evaluated


def test_fallbacks_keep_the_unfused_kernels(fake):
    from minpy.array_variants.cuda import cuda_core as cc
    from minpy_b200.lazy import LazyArray
    N, D, H = 64, 32, 16
    x, w, b = _arr((N, D)), _arr((D, H)), _arr((H,))
    gamma, beta, rm, rv = _arr((H,)), _arr((H,)), _arr((H,)), _arr((H,))
    # (1) an affine result consumed by something else evaluates the plain GEMM once
    a = cc.fused_affine(x, w, b)
    _ = a.ptr
    assert fake.names() == ['mc_gemm'] and fake.calls[0][1][12] == 0 and not a.pending()
    h = cc.fused_relu(a)                                   # concrete input: the unary kernel
    assert fake.names() == ['mc_gemm', 'mc_unary'] and h._tag is None          # pylint: disable=protected-access
    # (2) batchnorm whose output is read directly: the plain kernel, all three results bound
    del fake.calls[:]
    out, nrm, nrv = cc.fused_batchnorm(h, gamma, beta, rm, rv, 1e-5, 0.9)
    _ = nrm.ptr
    assert fake.names() == ['mc_batchnorm_train_fwd'] and not out.pending() and not nrv.pending()
    y = cc.fused_dropout(out, 0.5, 3, 0)                   # concrete input: the dropout kernel
    assert fake.names() == ['mc_batchnorm_train_fwd', 'mc_dropout'] and y._saved is None  # pylint: disable=protected-access
    # its VJPs: dropout launches, batchnorm without the relu / dropout peepholes
    del fake.calls[:]
    g = cc._dropout_vjp(y, out, 0.5, 3, 0)(_arr((N, H)))                       # pylint: disable=protected-access
    assert fake.names() == ['mc_dropout']
    dh, _, _ = cc._batchnorm_vjp((out, nrm, nrv), h, gamma, beta, rm, rv, 1e-5, 0.9)((g, None, None))  # pylint: disable=protected-access
    assert fake.names()[-1] == 'mc_bn_dropout_relu_bwd' and fake.calls[-1][1][11] == 0 and fake.calls[-1][1][15] == 0
    assert dh._tag is None                                                     # pylint: disable=protected-access
    # (3) shapes outside the fused batchnorm kernels: eager, un-fused
    del fake.calls[:]
    big = _arr((4096, 16))
    out2, _, _ = cc.fused_batchnorm(big, gamma, beta, rm, rv, 1e-5, 0.9)
    assert not isinstance(out2, LazyArray) and fake.names() == ['mc_batchnorm_train_fwd']
    # (4) MINPY_FUSE_FORWARD=0
    cc._FUSE_FORWARD, saved = False, cc._FUSE_FORWARD                          # pylint: disable=protected-access
    try:
        del fake.calls[:]
        a2 = cc.fused_affine(x, w, b)
        assert not isinstance(a2, LazyArray) and fake.names() == ['mc_gemm']
    finally:
        cc._FUSE_FORWARD = saved                                               # pylint: disable=protected-access


def test_dropout_marker_materialises_through_mc_dropout(fake):
    """A gradient recipe with a dropout marker that does NOT reach a batchnorm VJP is evaluated as
    chain segment -> mc_dropout -> chain segment."""
    from minpy_b200.lazy import ChainGrad, drop_step
    g = _arr((8, 8))
    other = _arr((8, 8))
    recipe = ChainGrad((8, 8), g).then(fake.MC_VJP_MUL_A, other).then(*drop_step(0.25, 11, 64)).then(fake.MC_VJP_NEG)
    recipe.materialize()
    assert fake.names() == ['mc_vjp_chain', 'mc_dropout', 'mc_vjp_chain']
    assert fake.calls[1][1][3:] == (0.25, 11, 64)
    del fake.calls[:]
    ChainGrad((8, 8), g).then(*drop_step(0.5, 1, 2)).materialize()
    assert fake.names() == ['mc_dropout'] and fake.calls[0][1][0] == g.ptr


def test_classifier_head_is_one_call(fake):
    """affine (<= 16 classes) -> softmax_loss: one mc_head_softmax_loss; the affine's data-gradient VJP returns the
    dh that kernel produced, its weight-gradient VJP still runs a GEMM on dlogits."""
    from minpy.array_variants.cuda import cuda_core as cc
    from minpy_b200.cuda_array import CudaArray, _I32
    from minpy_b200 import ops
    h, w, b = _arr((64, 32)), _arr((32, 10)), _arr((10,))
    lab = CudaArray.empty((64,), _I32)
    a = cc.fused_affine(h, w, b)
    loss = cc.fused_softmax_loss(a, lab)
    assert fake.names() == ['mc_head_softmax_loss'] and not a.pending()
    args = fake.calls[0][1]
    assert args[0] == h.ptr and args[1] == w.ptr and args[4] == a.ptr and args[-3:] == (64, 32, 10)
    dl = cc._softmax_loss_vjp(loss, a, lab)(1.0)                                # pylint: disable=protected-access
    dx_fn, dw_fn, _ = cc._affine_vjps()                                          # pylint: disable=protected-access
    dh = dx_fn(a, h, w, b)(dl)
    assert dh.ptr == args[6] and len(fake.calls) == 1
    dw_fn(a, h, w, b)(dl)
    ops.side_join()
    assert [n for n in fake.names() if not n.startswith('mc_side')] == ['mc_head_softmax_loss', 'mc_gemm']
    # more than 16 classes, or a concrete input: the plain loss kernel
    del fake.calls[:]
    cc.fused_softmax_loss(cc.fused_affine(h, _arr((32, 40)), _arr((40,))), lab)
    assert fake.names() == ['mc_gemm', 'mc_softmax_ce_fused']


def test_softmax_loss_gradient_comes_from_the_forward_kernel(fake):
    from minpy.array_variants.cuda import cuda_core as cc
    from minpy_b200.cuda_array import CudaArray, _I32
    x = _arr((32, 10))
    lab = CudaArray.empty((32,), _I32)
    loss = cc.fused_softmax_loss(x, lab)
    assert fake.names() == ['mc_softmax_ce_fused'] and fake.calls[0][1][3] is None          # no probabilities stored
    dx = cc._softmax_loss_vjp(loss, x, lab)(1.0)                                            # pylint: disable=protected-access
    assert dx.ptr == fake.calls[0][1][4] and len(fake.calls) == 1


def test_mlp_c3_step_launch_sequence(fake, monkeypatch):
    """One full `Solver._step` of BASELINE config 3 (5-layer MLP, batchnorm + dropout, batch 1024) against the
    recorder: the library calls it issues, through the real front end (layers, tape, optim).  Round 1 issued
    55 kernel launches per step; the fused step issues the 23 library calls below (a split-K GEMM is two launches)."""
    import minpy
    from minpy import array as marray
    from minpy.nn.io import DataBatch
    from minpy.nn.solver import Solver
    from minpy_b200 import workloads
    from minpy_b200.cuda_array import CudaArray, _I32
    minpy.set_global_policy('only_cuda')
    try:
        model = workloads.DeepMLP()
        solver = Solver(model, None, None, update_rule='sgd_momentum', optim_config={'learning_rate': 1e-3, 'momentum': 0.9},
                        init_rule='gaussian', init_config={'stdvar': 0.01}, verbose=False)
        solver.init()
        xb = marray.wrap(CudaArray.empty((1024, 3, 32, 32)))
        yb = marray.wrap(CudaArray.empty((1024,), _I32))
        solver._step(DataBatch([xb], [yb], pad=0, index=0))          # pylint: disable=protected-access
        del fake.calls[:]
        solver._step(DataBatch([xb], [yb], pad=0, index=1))          # pylint: disable=protected-access
        names = [n for n in fake.names() if n not in ('mc_readback_begin',)]
        # every layer's weight-gradient GEMM (4 hidden + the head) is bracketed onto the side stream; ONE join before the update
        # (5 weight gradients + the head's bias-gradient reduction; the hidden layers' bias gradients come out of
        # the fused batchnorm VJP)
        assert names.count('mc_side_begin') == 6 and names.count('mc_side_end') == 6 and names.count('mc_side_join') == 1
        for i, n in enumerate(names):
            if n == 'mc_side_begin':
                assert names[i + 1] in ('mc_gemm', 'mc_reduce') and names[i + 2] == 'mc_side_end'
        assert names.index('mc_side_join') == len(names) - 2 and names[-1] == 'mc_multi_update'
        kernels = [n for n in names if n not in ('mc_memset', 'mc_memcpy_d2d', 'mc_side_begin', 'mc_side_end', 'mc_side_join')]
        fwd_block = ['mc_gemm', 'mc_bn_dropout_fwd']
        bwd_block = ['mc_bn_dropout_relu_bwd', 'mc_gemm', 'mc_gemm']
        # (the head -- affine 512 -> 10, softmax_loss, both their chain VJPs -- is ONE call; its weight and bias
        # gradients are the 'mc_gemm', 'mc_reduce' pair on the side stream)
        want = fwd_block * 4 + ['mc_head_softmax_loss'] + \
            ['mc_gemm', 'mc_reduce'] + bwd_block * 3 + ['mc_bn_dropout_relu_bwd', 'mc_gemm'] + ['mc_multi_update']
        assert sorted(kernels) == sorted(want), kernels
        assert kernels[:9] == want[:9]                                # the forward order
        assert kernels.count('mc_gemm') == 12 and len(kernels) == 23
        relu_modes = [c[1][12] for c in fake.calls if c[0] == 'mc_gemm'][:4]
        assert relu_modes == [2, 2, 2, 2]
        bwd = [c[1] for c in fake.calls if c[0] == 'mc_bn_dropout_relu_bwd']
        assert all(a[11] == 1 and a[15] == 1 and a[8] is not None for a in bwd)     # dropout + relu + bias gradient fused
    finally:
        minpy.set_global_policy('prefer_cuda')


def test_side_stream_only_for_sections_that_leave_room(fake):
    """ops.side_stream: operands beyond SIDE_MAX_BYTES (a contraction that fills the GPU by itself, e.g. the LSTM's
    weight gradients over 210 MB of dA) run on the compute stream; small sections and sections marked any_size (the
    convolution weight gradients, measured) go to the side stream."""
    from minpy_b200 import ops
    small, big = _arr((1024, 512)), _arr((12800, 4096))
    assert big.nbytes > ops.SIDE_MAX_BYTES > small.nbytes
    with ops.side_stream(small) as on_side:
        assert on_side
    assert fake.names() == ['mc_side_begin', 'mc_side_end']
    ops.side_join()
    del fake.calls[:]
    with ops.side_stream(small, big) as on_side:
        assert not on_side
    assert fake.names() == []
    with ops.side_stream(small, big, any_size=True) as on_side:
        assert on_side
    assert fake.names() == ['mc_side_begin', 'mc_side_end']
    ops.side_join()
    assert fake.names()[-1] == 'mc_side_join'


def test_transpose_into_a_given_buffer(fake):
    """ops.transpose(out=): one strided copy into the caller's array (the LSTM VJP writes the time-major gradient
    straight into dH's slots 1..T); the shape is checked."""
    from minpy_b200 import ops
    x, out = _arr((4, 6, 8)), _arr((6, 4, 8))
    assert ops.transpose(x, (1, 0, 2), out=out) is out
    assert fake.names() == ['mc_strided_copy'] and fake.calls[0][1][1] == out.ptr
    with pytest.raises(ValueError):
        ops.transpose(x, (1, 0, 2), out=_arr((4, 6, 8)))
