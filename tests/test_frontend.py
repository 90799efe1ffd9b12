"""GPU parity of the MinPy front end on the `cuda` array variant against the NumPy oracle.

Everything here goes through the user-facing API (`import minpy.numpy as np`, operators on
Arrays, `core.grad_and_loss`, `nn.layers`, `nn.optim`, `nn.solver.Solver`) under the
`only_cuda` policy, i.e. through the registry / Primitive / rewritten tape and then the C ABI
of libminpy_cuda.so.  The oracle (oracle/minpy_oracle.py) is the checker only.

Tolerances (north_star): fp32 elementwise / reduction 1e-5 relative; TF32 tensor-core GEMM
paths 1e-2 relative; argmax / indexing / label gather bit-exact.
"""
import numpy as np
import pytest

ELEM_TOL = 1e-5
GEMM_TOL = 1e-2
KINK_TOL = 5e-2


def rel(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape, (a.shape, b.shape)
    if a.size == 0:
        return 0.0
    return float(np.max(np.abs(a - b)) / max(1e-30, np.max(np.abs(b))))


def rel_l2(a, b):
    """Relative Frobenius error.  Used where a relu sits behind a TF32 GEMM: a pre-activation
    within TF32 rounding (~1e-3 absolute at these sizes) of zero lands on the other side of the
    kink than in the float64 oracle for ~0.05% of the units, and each flip moves one column of
    dW by one sample's whole contribution.  That is a property of relu under any reduced
    precision, not of the kernel (the GEMM itself is checked at 1e-2 in test_kernels_gpu.py),
    so those nets are checked at KINK_TOL in Frobenius norm under TF32 and in max-norm at 1e-4
    in the exact-accumulate FP32 GEMM mode."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape, (a.shape, b.shape)
    return float(np.linalg.norm(a - b) / max(1e-30, np.linalg.norm(b)))


@pytest.fixture(scope='module', params=[pytest.param('only_cuda', marks=pytest.mark.gpu), 'only_numpy'])
def mp(request):
    """(minpy.numpy namespace, core, layers, optim, oracle) under one policy: `only_cuda` is the
    product path (GPU box); `only_numpy` runs the same front end on the host variant, which
    checks the Array / Primitive / tape logic and the oracle against each other on CPU."""
    import minpy
    minpy.set_global_policy(request.param)
    import minpy.numpy as mnp
    from minpy import core
    from minpy.nn import layers, optim
    from oracle import minpy_oracle as orc
    yield mnp, core, layers, optim, orc
    minpy.set_global_policy('prefer_cuda')


def f32(rng, *shape):
    return rng.randn(*shape).astype(np.float32)


# ---------------------------------------------------------------- primitives + VJP chains --
UNARY_CASES = ['negative', 'abs', 'sqrt', 'exp', 'log', 'sin', 'cos', 'tanh']


@pytest.mark.parametrize('name', UNARY_CASES)
def test_unary_forward_and_grad(mp, name):
    mnp, core, _, _, orc = mp
    rng = np.random.RandomState(1)
    x = np.abs(f32(rng, 37, 19)) + 0.5
    w = f32(rng, 37, 19)

    def fn(ns):
        return lambda a: ns.sum(getattr(ns, name)(a) * w)
    g, val = core.grad_and_loss(fn(mnp))(mnp.array(x))
    og, oval = orc.grad_and_loss(fn(orc))(x.astype(np.float64))
    assert rel(val.asnumpy(), oval) <= ELEM_TOL
    assert rel(g.asnumpy(), og) <= ELEM_TOL


@pytest.mark.parametrize('op', ['add', 'subtract', 'multiply', 'divide', 'maximum', 'minimum', 'power'])
@pytest.mark.parametrize('shapes', [((8, 12), (8, 12)), ((8, 12), (12,)), ((8, 12), (8, 1)), ((5, 8, 12), (8, 1))])
def test_binary_broadcast_grads(mp, op, shapes):
    mnp, core, _, _, orc = mp
    rng = np.random.RandomState(2)
    a = np.abs(f32(rng, *shapes[0])) + 0.5
    b = np.abs(f32(rng, *shapes[1])) + 0.5
    w = f32(rng, *np.broadcast_shapes(*shapes))

    def fn(ns):
        return lambda p, q: ns.sum(getattr(ns, op)(p, q) * w)
    (ga, gb), val = core.grad_and_loss(fn(mnp), argnum=[0, 1])(mnp.array(a), mnp.array(b))
    (oa, ob), oval = orc.grad_and_loss(fn(orc), argnum=[0, 1])(a.astype(np.float64), b.astype(np.float64))
    assert rel(val.asnumpy(), oval) <= 1e-4
    assert rel(ga.asnumpy(), oa) <= 1e-4     # `power` goes through powf/logf: 1e-4
    assert rel(gb.asnumpy(), ob) <= 1e-4


def test_operator_chain_fuses_and_matches(mp):
    """A chain of elementwise ops: the backward pass must match the oracle and launch far
    fewer kernels than one per VJP (the tape rewrite)."""
    import ctypes
    import minpy
    from minpy.dispatch.policy import OnlyCudaPolicy
    if not isinstance(minpy.get_global_policy(), OnlyCudaPolicy):
        pytest.skip('launch counting is about the cuda variant')
    from minpy_b200._lib import get_lib
    mnp, core, _, _, orc = mp
    rng = np.random.RandomState(3)
    x = f32(rng, 256, 128)
    y = np.abs(f32(rng, 256, 128)) + 1.0

    def fn(ns):
        def f(a):
            h = ns.tanh(a * 0.5 + 1.0)
            h = ns.exp(-h) / y
            h = ns.sqrt(h * h + 1.0)
            return ns.sum(h)
        return f
    lib = get_lib()
    n0, n1, n2 = ctypes.c_uint64(), ctypes.c_uint64(), ctypes.c_uint64()
    lib.mc_launch_count(ctypes.byref(n0))
    g, val = core.grad_and_loss(fn(mnp))(mnp.array(x))
    lib.mc_launch_count(ctypes.byref(n1))
    og, oval = orc.grad_and_loss(fn(orc))(x.astype(np.float64))
    assert rel(val.asnumpy(), oval) <= ELEM_TOL
    assert rel(g.asnumpy(), og) <= 5e-5
    # forward: 9 elementwise kernels + 1 sum; backward must be a handful of chain launches
    # (`h * h` fans out, which forces one intermediate), not one launch per VJP record (>= 11)
    lib.mc_launch_count(ctypes.byref(n2))
    assert n1.value - n0.value <= 10 + 6, n1.value - n0.value


@pytest.mark.parametrize('axis', [None, 0, 1])
@pytest.mark.parametrize('keepdims', [False, True])
@pytest.mark.parametrize('red', ['sum', 'max', 'min'])
def test_reduction_known_answers(mp, red, axis, keepdims):
    """The exact integer-valued gradients the reference pins in
    tests/unittest/test_autograd.py:89-160 (ties receive the full gradient each)."""
    mnp, core, _, _, _ = mp
    x = np.array([[1., 2., 2.], [0., 2., -1.]], dtype=np.float32)
    fn = lambda a: getattr(mnp, red)(a, axis=axis, keepdims=keepdims)
    g = core.grad(fn)(mnp.array(x)).asnumpy()
    if red == 'sum':
        want = np.ones_like(x)
    else:
        ext = getattr(np, red)(x, axis=axis, keepdims=True)
        want = (x == ext).astype(np.float32)
    assert np.array_equal(g, want), (g, want)


def test_argmax_and_compare_bit_exact(mp):
    mnp, _, _, _, _ = mp
    rng = np.random.RandomState(4)
    x = f32(rng, 513, 10)
    x[7, 3] = x[7, 8] = 9.0      # a tie: first index wins
    lab = rng.randint(0, 10, 513)
    am = mnp.argmax(mnp.array(x), axis=1)
    assert np.array_equal(am.asnumpy(), np.argmax(x, axis=1))
    correct = mnp.sum(am == mnp.array(lab))
    assert int(correct.asnumpy()) == int(np.sum(np.argmax(x, axis=1) == lab))


def test_getitem_slices_and_gather_grads(mp):
    mnp, core, _, _, orc = mp
    rng = np.random.RandomState(5)
    x = f32(rng, 16, 24)
    idx = rng.randint(0, 24, 16)

    def fn(ns, index_of):
        def f(a):
            s1 = a[:, 0:8]
            s2 = a[:, 8:16]
            s3 = a[:, ::2]
            picked = a[ns.arange(16), index_of]
            return ns.sum(s1 * s2) + ns.sum(s3) * 2.0 + ns.sum(picked * picked)
        return f
    g, val = core.grad_and_loss(fn(mnp, mnp.array(idx)))(mnp.array(x))
    og, oval = orc.grad_and_loss(fn(orc, idx))(x.astype(np.float64))
    assert rel(val.asnumpy(), oval) <= ELEM_TOL
    assert rel(g.asnumpy(), og) <= ELEM_TOL
    # forward gathers are bit-exact
    a = mnp.array(x)
    assert np.array_equal(a[:, 3:11].asnumpy(), x[:, 3:11])
    assert np.array_equal(a[mnp.arange(16), mnp.array(idx)].asnumpy(), x[np.arange(16), idx])


# ------------------------------------------------------------------------------ layers ----
def test_affine_relu_softmax_loss(mp):
    mnp, core, layers, _, orc = mp
    rng = np.random.RandomState(6)
    N, D, H, C = 128, 192, 256, 10
    x, y = f32(rng, N, D), rng.randint(0, C, N)
    ps = [f32(rng, D, H) * 0.1, f32(rng, H) * 0.1, f32(rng, H, C) * 0.1, f32(rng, C) * 0.1]

    def fn(L, xx, yy):
        def f(w1, b1, w2, b2):
            return L.softmax_loss(L.affine(L.relu(L.affine(xx, w1, b1)), w2, b2), yy)
        return f
    ograds, oloss = orc.grad_and_loss(fn(orc, x.astype(np.float64), y), argnum=range(4))(
        *[p.astype(np.float64) for p in ps])
    grads, loss = core.grad_and_loss(fn(layers, mnp.array(x), mnp.array(y)), argnum=range(4))(
        *[mnp.array(p) for p in ps])
    assert rel(loss.asnumpy(), oloss) <= GEMM_TOL
    for g, og in zip(grads, ograds):
        assert rel_l2(g.asnumpy(), og) <= KINK_TOL        # TF32 tensor-core GEMMs + relu kinks
    from minpy_b200 import ops
    ops.set_gemm_precision('fp32')                         # FP32-accumulate cross-check
    try:
        grads, loss = core.grad_and_loss(fn(layers, mnp.array(x), mnp.array(y)), argnum=range(4))(
            *[mnp.array(p) for p in ps])
    finally:
        ops.set_gemm_precision('tf32')
    assert rel(loss.asnumpy(), oloss) <= 1e-5
    for g, og in zip(grads, ograds):
        assert rel(g.asnumpy(), og) <= 1e-4


def test_unfused_layer_compositions_match_fused(mp):
    """The same net written with raw primitives (np.dot + b, np.maximum(0, .), softmax_output
    composition) -- the reference's own formulation -- gives the oracle's numbers too."""
    mnp, core, _, _, orc = mp
    rng = np.random.RandomState(7)
    N, D, C = 64, 48, 10
    x, y = f32(rng, N, D), rng.randint(0, C, N)
    w, b = f32(rng, D, C) * 0.1, f32(rng, C) * 0.1

    def fn(ns, xx, yy):
        def f(w_, b_):
            s = ns.maximum(0, ns.dot(xx, w_) + b_)
            onehot = ns.zeros([N, C])
            ns.onehot_encode(yy, onehot)
            prob = ns.softmax_output(s, onehot)
            return -ns.sum(ns.log(prob) * onehot) / N
        return f
    (gw, gb), loss = core.grad_and_loss(fn(mnp, mnp.array(x), mnp.array(y)), argnum=[0, 1])(mnp.array(w), mnp.array(b))
    (ow, ob), oloss = orc.grad_and_loss(fn(orc, x.astype(np.float64), y), argnum=[0, 1])(
        w.astype(np.float64), b.astype(np.float64))
    assert rel(loss.asnumpy(), oloss) <= GEMM_TOL
    assert rel(gw.asnumpy(), ow) <= GEMM_TOL
    assert rel(gb.asnumpy(), ob) <= GEMM_TOL


def test_batchnorm_train_and_test(mp):
    mnp, core, layers, _, orc = mp
    rng = np.random.RandomState(8)
    N, D = 256, 96
    x = f32(rng, N, D) * 2.0 + 0.5
    gamma, beta = np.abs(f32(rng, D)) + 0.5, f32(rng, D)
    rm, rv = f32(rng, D) * 0.1, np.abs(f32(rng, D)) + 0.5
    w = f32(rng, N, D)

    def fn(L, ns, rm_, rv_):
        def f(x_, g_, b_):
            out, _, _ = L.batchnorm(x_, g_, b_, running_mean=rm_, running_var=rv_)
            return ns.sum(out * w)
        return f
    grads, val = core.grad_and_loss(fn(layers, mnp, mnp.array(rm), mnp.array(rv)), argnum=[0, 1, 2])(
        mnp.array(x), mnp.array(gamma), mnp.array(beta))
    ograds, oval = orc.grad_and_loss(fn(orc, orc, rm.astype(np.float64), rv.astype(np.float64)), argnum=[0, 1, 2])(
        x.astype(np.float64), gamma.astype(np.float64), beta.astype(np.float64))
    assert rel(val.asnumpy(), oval) <= 1e-4
    for g, og in zip(grads, ograds):
        assert rel(g.asnumpy(), og) <= 1e-4
    out, nrm, nrv = layers.batchnorm(mnp.array(x), mnp.array(gamma), mnp.array(beta), running_mean=mnp.array(rm),
                                     running_var=mnp.array(rv))
    oout, orm, orv = orc.batchnorm(x.astype(np.float64), gamma.astype(np.float64), beta.astype(np.float64),
                                   running_mean=rm.astype(np.float64), running_var=rv.astype(np.float64))
    assert rel(out.asnumpy(), orc.raw(oout)) <= ELEM_TOL * 5
    assert rel(nrm.asnumpy(), orc.raw(orm)) <= ELEM_TOL * 5
    assert rel(nrv.asnumpy(), orc.raw(orv)) <= ELEM_TOL * 5
    tout, _, _ = layers.batchnorm(mnp.array(x), mnp.array(gamma), mnp.array(beta), mode='test',
                                  running_mean=mnp.array(rm), running_var=mnp.array(rv))
    otout, _, _ = orc.batchnorm(x.astype(np.float64), gamma.astype(np.float64), beta.astype(np.float64), mode='test',
                                running_mean=rm.astype(np.float64), running_var=rv.astype(np.float64))
    assert rel(tout.asnumpy(), orc.raw(otout)) <= ELEM_TOL * 5


def test_dropout_semantics(mp):
    """Not inverted (survivors unscaled), test mode scales by 1-p, the backward mask is the
    forward mask (nn/layers.py:121-147)."""
    mnp, core, layers, _, _ = mp
    rng = np.random.RandomState(9)
    x = np.abs(f32(rng, 512, 256)) + 0.1
    a = mnp.array(x)
    out = layers.dropout(a, 0.3).asnumpy()
    kept = out != 0
    assert np.array_equal(out[kept], x[kept])
    assert abs(kept.mean() - 0.7) < 0.01
    assert rel(layers.dropout(a, 0.3, mode='test').asnumpy(), x * 0.7) <= ELEM_TOL
    holder = {}

    def f(a_):
        holder['out'] = layers.dropout(a_, 0.5)
        return mnp.sum(holder['out'] * 3.0)
    g = core.grad(f)(a).asnumpy()
    mask = holder['out'].asnumpy() != 0
    assert np.array_equal(g, np.where(mask, np.float32(3.0), np.float32(0.0)))
    o1 = layers.dropout(a, 0.5, seed=11).asnumpy()
    o2 = layers.dropout(a, 0.5, seed=11).asnumpy()
    assert np.array_equal(o1, o2)


def test_lstm_step_fused_vs_oracle(mp):
    mnp, core, layers, _, orc = mp
    rng = np.random.RandomState(10)
    N, D, H = 32, 48, 64
    x, h, c = f32(rng, N, D), f32(rng, N, H), f32(rng, N, H)
    Wx, Wh, b = f32(rng, D, 4 * H) * 0.2, f32(rng, H, 4 * H) * 0.2, f32(rng, 4 * H) * 0.2
    wh, wc = f32(rng, N, H), f32(rng, N, H)

    def fn(L, ns):
        def f(*args):
            nh, nc = L.lstm_step(*args)
            return ns.sum(nh * wh) + ns.sum(nc * wc)
        return f
    args = [x, h, c, Wx, Wh, b]
    grads, val = core.grad_and_loss(fn(layers, mnp), argnum=range(6))(*[mnp.array(a) for a in args])
    ograds, oval = orc.grad_and_loss(fn(orc, orc), argnum=range(6))(*[a.astype(np.float64) for a in args])
    assert rel(val.asnumpy(), oval) <= GEMM_TOL
    for g, og in zip(grads, ograds):
        assert rel(g.asnumpy(), og) <= GEMM_TOL


def test_lstm_temporal_training_loss_path(mp):
    """The examples/nn/lstm.py shape of computation at a small size: slicing x[:, t, :], lstm
    steps, np.append of hidden states, temporal affine and masked temporal softmax loss."""
    mnp, core, layers, _, orc = mp
    rng = np.random.RandomState(11)
    N, T, V, H = 8, 6, 12, 16
    tok = rng.randint(0, V, (N, T))
    x = np.eye(V, dtype=np.float32)[tok]
    y = rng.randint(0, V, (N, T))
    mask = np.ones((N, T), np.float32)
    ps = [f32(rng, V, 4 * H) * 0.3, f32(rng, H, 4 * H) * 0.3, f32(rng, 4 * H) * 0.1, f32(rng, H, V) * 0.3,
          f32(rng, V) * 0.1]

    def fn(L, ns, xx, yy, mm, h0):
        def f(Wx, Wh, b, Wa, ba):
            hs = L.lstm_temporal(xx, h0, Wx, Wh, b) if hasattr(L, 'lstm_temporal') else None
            return L.temporal_softmax_loss(L.temporal_affine(hs, Wa, ba), yy, mm)
        return f

    def oracle_fn(xx, yy, mm):
        def f(Wx, Wh, b, Wa, ba):
            ph, pc = np.zeros((N, H)), np.zeros((N, H))
            hs = []
            for t in range(T):
                ph, pc = orc.lstm_step(xx[:, t, :], ph, pc, Wx, Wh, b)
                hs.append(ph.reshape(N, 1, H))
            hcat = hs[0]
            for nxt in hs[1:]:
                hcat = orc.append(hcat, nxt, axis=1)
            return orc.temporal_softmax_loss(orc.temporal_affine(hcat, Wa, ba), yy, mm)
        return f
    grads, val = core.grad_and_loss(
        fn(layers, mnp, mnp.array(x), mnp.array(y), mnp.array(mask), mnp.zeros((N, H))), argnum=range(5))(
            *[mnp.array(p) for p in ps])
    ograds, oval = orc.grad_and_loss(oracle_fn(x.astype(np.float64), y, mask.astype(np.float64)), argnum=range(5))(
        *[p.astype(np.float64) for p in ps])
    assert rel(val.asnumpy(), oval) <= GEMM_TOL
    for g, og in zip(grads, ograds):
        assert rel(g.asnumpy(), og) <= GEMM_TOL


@pytest.mark.parametrize('rule', ['sgd', 'sgd_momentum', 'rmsprop', 'adam'])
def test_update_rules_two_steps(mp, rule):
    mnp, _, _, optim, orc = mp
    rng = np.random.RandomState(12)
    w, d1, d2 = f32(rng, 300, 7), f32(rng, 300, 7), f32(rng, 300, 7)
    cw, cfg = mnp.array(w), {'learning_rate': 1e-2}
    ow, ocfg = w.astype(np.float64), {'learning_rate': 1e-2}
    for d in (d1, d2):
        cw, cfg = getattr(optim, rule)(cw, mnp.array(d), cfg)
        ow, ocfg = getattr(orc, rule)(ow, d.astype(np.float64), ocfg)
    assert rel(cw.asnumpy(), ow) <= ELEM_TOL


# ------------------------------------------------------------------------------ solver ----
def _mlp_model(hidden, classes, in_dim, bn_dropout=False):
    from minpy.nn import layers
    from minpy.nn.model import ModelBase

    class Net(ModelBase):
        def __init__(self):
            super(Net, self).__init__()
            dims = [in_dim] + list(hidden) + [classes]
            self.n = len(dims) - 1
            for i in range(self.n):
                self.add_param('w%d' % i, (dims[i], dims[i + 1])).add_param('b%d' % i, (dims[i + 1],))

        def forward(self, X, mode):
            h = X.reshape((X.shape[0], in_dim)) if len(X.shape) != 2 else X
            for i in range(self.n - 1):
                h = layers.relu(layers.affine(h, self.params['w%d' % i], self.params['b%d' % i]))
            return layers.affine(h, self.params['w%d' % (self.n - 1)], self.params['b%d' % (self.n - 1)])

        def loss(self, predict, y):
            return layers.softmax_loss(predict, y)
    return Net()


def test_solver_steps_match_oracle_training(mp):
    """Config C1 in miniature (MLP, sgd_momentum) for 5 steps through Solver._step: the loss
    curve and the final parameters follow the oracle's training run."""
    mnp, _, _, _, orc = mp
    from minpy.nn.io import NDArrayIter
    from minpy.nn.solver import Solver
    rng = np.random.RandomState(13)
    B, D, H, C, steps = 64, 3 * 8 * 8, 96, 10, 5
    X = f32(rng, B * steps, 3, 8, 8)
    Y = rng.randint(0, C, B * steps)
    init = {'w0': f32(rng, D, H) * 0.05, 'b0': np.zeros(H, np.float32), 'w1': f32(rng, H, C) * 0.05,
            'b1': np.zeros(C, np.float32)}
    model = _mlp_model([H], C, D)
    it = NDArrayIter(X, Y, batch_size=B, shuffle=False)
    solver = Solver(model, it, it, update_rule='sgd_momentum', optim_config={'learning_rate': 1e-2, 'momentum': 0.9},
                    init_rule='gaussian', verbose=False)
    solver.init()
    for k, v in init.items():
        model.params[k] = mnp.array(v)
    for batch in it:
        solver._step(batch)  # pylint: disable=protected-access

    params = dict((k, v.astype(np.float64)) for k, v in init.items())
    cfgs = dict((k, {'learning_rate': 1e-2, 'momentum': 0.9}) for k in params)
    names = ['w0', 'b0', 'w1', 'b1']
    olosses = []
    for s in range(steps):
        xb = X[s * B:(s + 1) * B].reshape(B, D).astype(np.float64)
        yb = Y[s * B:(s + 1) * B]

        def f(w0, b0, w1, b1):
            return orc.softmax_loss(orc.affine(orc.relu(orc.affine(xb, w0, b0)), w1, b1), yb)
        grads, loss = orc.grad_and_loss(f, argnum=range(4))(*[params[n] for n in names])
        olosses.append(float(loss))
        for n, g in zip(names, grads):
            params[n], cfgs[n] = orc.sgd_momentum(params[n], g, cfgs[n])
    got = [float(l) for l in solver.loss_history]
    assert np.allclose(got, olosses, rtol=GEMM_TOL), (got, olosses)
    for n in names:
        assert rel(model.params[n].asnumpy(), params[n]) <= GEMM_TOL


def test_solver_check_accuracy_on_device(mp):
    mnp, _, _, _, _ = mp
    from minpy.nn.io import NDArrayIter
    from minpy.nn.solver import Solver
    rng = np.random.RandomState(14)
    B, D, C = 50, 32, 5
    X = f32(rng, 200, D)
    W = f32(rng, D, C)
    Y = np.argmax(X.dot(W), axis=1)
    model = _mlp_model([], C, D)
    it = NDArrayIter(X, Y, batch_size=B)
    solver = Solver(model, it, it, verbose=False)
    solver.init()
    model.params['w0'] = mnp.array(W)
    model.params['b0'] = mnp.zeros(C)
    assert solver.check_accuracy(it) == 1.0


@pytest.mark.gpu
@pytest.mark.parametrize('rule', ['sgd', 'sgd_momentum', 'rmsprop', 'adam'])
def test_multi_tensor_update_equals_per_parameter_loop(rule):
    """optim.multi_update (one launch for all parameters, used by Solver._step) must leave the
    same parameters AND the same config contents as the reference's per-parameter loop
    (nn/solver.py:190-195), step after step, including tensors of odd sizes."""
    import minpy
    minpy.set_global_policy('only_cuda')
    import minpy.numpy as mnp
    from minpy.nn import optim
    rng = np.random.RandomState(12)
    shapes = [(33, 17), (5,), (128, 64), (1,), (7, 3, 3, 3), (300, 70)] * 7   # 42 tensors: more than one launch group; whole (float4 path) and ragged tiles
    w0 = [rng.randn(*s).astype(np.float32) for s in shapes]
    fn = getattr(optim, rule)
    base = {'learning_rate': 0.05}
    pa, pb = [mnp.array(w) for w in w0], [mnp.array(w) for w in w0]
    ca, cb = [dict(base) for _ in shapes], [dict(base) for _ in shapes]
    for step in range(3):
        grads = [rng.randn(*s).astype(np.float32) for s in shapes]
        ga, gb = [mnp.array(g) for g in grads], [mnp.array(g) for g in grads]
        out = optim.multi_update(fn, pa, ga, ca)
        assert out is not None, 'fast path must apply to device tensors'
        pa, ca = out
        for i in range(len(shapes)):
            pb[i], cb[i] = fn(pb[i], gb[i], cb[i])
        for i in range(len(shapes)):
            assert pa[i].shape == pb[i].shape
            # same fp32 formulas; the two kernels may contract a*b+c into an fma differently: 1 ulp of
            # the update term, which after cancellation in w - update can be several ulp of a small
            # w (seen: 1.01e-6 on a 1-element tensor) -> the north_star fp32 elementwise bound, 1e-5
            assert rel(pa[i].asnumpy(), pb[i].asnumpy()) <= 1e-5, (rule, step, i)
            assert set(ca[i].keys()) == set(cb[i].keys())
            for k in ca[i]:
                va, vb = ca[i][k], cb[i][k]
                if hasattr(va, 'asnumpy'):
                    assert rel(va.asnumpy(), vb.asnumpy()) <= 1e-5, (rule, step, i, k)
                else:
                    assert va == vb, (rule, k)
    # host arrays -> no fast path
    assert optim.multi_update(fn, [w0[0]], [w0[0]], [{}]) is None
    minpy.set_global_policy('prefer_cuda')
