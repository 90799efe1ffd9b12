"""Scope-table rows that were "one test away" (VERDICT r1 #8): a7 `Context` scoping, a28
`quick_grad_check`, `primitive.customop`, f3 checkpoint / resume with optimizer state, f4 the
`prefer_cuda` fallback to NumPy for names without a kernel + op statistics.

Each has a CPU case on the NumPy variant of the re-created front end and a GPU case on the cuda
variant.  Reference: minpy/context.py:92-153, minpy/utils/gradient_checker.py:11-62,
minpy/primitive.py:462-483 + tests/unittest/test_customop.py:14-77, minpy/nn/model.py:141-163,
minpy/dispatch/policy.py:36-131,220-231 + docs/feature/op_stat.rst.
"""
import io
import os
import sys
from contextlib import redirect_stdout

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


@pytest.fixture(params=[pytest.param('only_cuda', marks=pytest.mark.gpu), 'only_numpy'])
def policy(request):
    import minpy
    minpy.set_global_policy(request.param)
    yield request.param
    minpy.set_global_policy('prefer_cuda')


# ------------------------------------------------------------------------------ a7 --------
def test_context_scoping_and_set_context():
    """`with gpu(i):` scopes the default context and restores it; set_context round-trips."""
    from minpy import context
    from minpy.context import Context, cpu, gpu, set_context, current_context
    start = current_context()
    assert start == cpu(0) and str(start) == 'cpu(0)'
    with gpu(1) as c:
        assert current_context() is c and c.cuda_device() == 1 and str(c) == 'gpu(1)'
        with cpu(0):
            assert current_context() == cpu(0) and current_context().cuda_device() is None
        assert current_context() == gpu(1)
    assert current_context() == start
    set_context(gpu(0))
    try:
        assert Context.default_ctx == gpu(0) and context.current_context().device_type == 'gpu'
    finally:
        set_context(start)
    assert current_context() == cpu(0)
    with pytest.raises(ValueError):
        Context('tpu', 0)
    assert hash(gpu(3)) == hash(Context('gpu', 3)) and gpu(3) != gpu(2)


@pytest.mark.gpu
def test_arrays_record_their_context_on_one_gpu():
    import minpy
    minpy.set_global_policy('only_cuda')
    import minpy.numpy as mnp
    from minpy.array_variants import ArrayType
    from minpy.context import gpu
    try:
        with gpu(0):
            x = mnp.ones((4, 4)) * 3.0
            assert x.context == gpu(0)
        raw = x.get_data(ArrayType.CUDA)
        assert raw.device == 0 and raw.context == ('gpu', 0)
        assert np.all(x.asnumpy() == 3.0)
    finally:
        minpy.set_global_policy('prefer_cuda')


@pytest.mark.gpu
def test_with_gpu1_places_primitives_on_device_1():
    """2-GPU box: a primitive inside `with gpu(1):` runs on device 1 and its result lives there;
    operands from device 0 are copied over transiently; outside the scope work is on device 0 again;
    gradients of a function evaluated on device 1 are computed there even when `grad` returns outside."""
    import ctypes
    import minpy
    from minpy_b200._lib import get_lib
    count = ctypes.c_int()
    get_lib().mc_device_count(ctypes.byref(count))
    if count.value < 2:
        pytest.skip('needs 2 GPUs')
    minpy.set_global_policy('only_cuda')
    import minpy.numpy as mnp
    from minpy import core
    from minpy.array_variants import ArrayType
    from minpy.context import gpu
    try:
        rng = np.random.RandomState(0)
        a, b = rng.randn(64, 96).astype(np.float32), rng.randn(96, 32).astype(np.float32)
        x0 = mnp.array(a)
        assert x0.get_data(ArrayType.CUDA).device == 0
        with gpu(1):
            w1 = mnp.array(b)
            y1 = mnp.dot(x0, w1) + 1.0                 # x0 lives on device 0: copied over for this call
            assert w1.get_data(ArrayType.CUDA).device == 1 and y1.get_data(ArrayType.CUDA).device == 1
            g = core.grad(lambda w: mnp.sum(mnp.dot(x0, w) * 2.0))(w1)
            assert g.get_data(ArrayType.CUDA).device == 1
        z0 = x0 * 2.0                                  # back on the home device
        assert z0.get_data(ArrayType.CUDA).device == 0
        assert x0.get_data(ArrayType.CUDA).device == 0      # the operand itself never moved
        np.testing.assert_allclose(y1.asnumpy(), a.dot(b) + 1.0, rtol=1e-2, atol=1e-2)
        np.testing.assert_allclose(g.asnumpy(), 2.0 * a.T.dot(np.ones((64, 32), np.float32)), rtol=1e-2, atol=1e-2)
        np.testing.assert_allclose(z0.asnumpy(), 2.0 * a, rtol=1e-6)
    finally:
        minpy.set_global_policy('prefer_cuda')


# ------------------------------------------------------------------------------ a28 -------
def test_quick_grad_check(policy):
    """gradient_checker.quick_grad_check on a function of the bound variant; and it must FAIL for a
    deliberately wrong gradient (a custom primitive whose VJP is off by a factor)."""
    import minpy.numpy as mnp
    from minpy.primitive import customop
    from minpy.utils.gradient_checker import quick_grad_check
    rng = np.random.RandomState(3)
    w = rng.randn(12, 5)
    x = mnp.array(rng.randn(7, 12))

    def loss(wv):
        h = mnp.tanh(mnp.dot(x, wv))
        return mnp.sum(h * h) / 7.0
    assert quick_grad_check(loss, mnp.array(w), verbose=False, rs=np.random.RandomState(1))

    @customop('numpy')
    def cube(v):
        return v ** 3
    cube.def_grad(lambda ans, v: lambda g: g * 2.0 * v ** 2)        # wrong on purpose: should be 3 v^2
    assert not quick_grad_check(lambda v: mnp.sum(cube(v)), mnp.array(rng.randn(6) + 2.0), verbose=False,
                                rs=np.random.RandomState(2))


# -------------------------------------------------------------------------- customop ------
_X = np.array([[0.24854138, 1.94385293, 2.33848549, 2.75407309, 1.66905118],
               [0.26498274, 1.60618255, 1.25387436, 2.9215846, 1.26427169],
               [0.87108803, 1.45227827, 2.17339809, 0.50049702, 0.5883466],
               [0.66406034, 0.48855862, 1.53960508, 0.66568797, 2.23948055],
               [2.72220612, 1.82959485, 1.51552618, 1.54757016, 1.64023012],
               [1.69430802, 2.21234513, 0.44159807, 1.94465274, 0.11623679],
               [0.71774937, 1.99183721, 2.93154152, 0.23254174, 1.63623933],
               [1.54450952, 2.32885258, 1.64220968, 1.66349828, 2.50975782],
               [0.99172053, 2.60171951, 1.14377575, 0.28264201, 2.50368237],
               [1.99669231, 2.16996937, 1.77290071, 1.34783694, 2.42391734]])
_LABEL = np.array([4, 0, 0, 1, 0, 4, 0, 2, 1, 3])
_LOSS = 2.16612911224                 # tests/unittest/test_customop.py:62
_GRAD_ROW0 = np.array([0.00323393, 0.01761957, 0.02614461, 0.0396159, -0.08661401])   # :64


def _softmax_numpy():
    from minpy.primitive import customop

    @customop('numpy')
    def softmax(x, y):
        y = y.astype(int)
        probs = np.exp(x - np.max(x, axis=1, keepdims=True))
        probs /= np.sum(probs, axis=1, keepdims=True)
        return -np.sum(np.log(probs[np.arange(x.shape[0]), y])) / x.shape[0]

    def softmax_grad(ans, x, label):
        def grad(g):
            y = label.astype(int)
            probs = np.exp(x - np.max(x, axis=1, keepdims=True))
            probs /= np.sum(probs, axis=1, keepdims=True)
            dx = probs.copy()
            dx[np.arange(x.shape[0]), y] -= 1
            return dx / x.shape[0]
        return grad
    softmax.def_grad(softmax_grad)
    return softmax


def test_customop_numpy_reference_literals():
    """tests/unittest/test_customop.py:14-77, on the NumPy variant (float64; the literals were printed from float32: 1e-7)."""
    import minpy
    minpy.set_global_policy('only_numpy')
    try:
        import minpy.numpy as mnp
        from minpy.core import grad
        softmax = _softmax_numpy()
        x, label = mnp.array(_X), mnp.array(_LABEL)
        assert abs(float(softmax(x, label)) - _LOSS) < 1e-7      # the literal went through float32
        g = grad(softmax)(x, label).asnumpy()
        assert np.max(np.abs(g[0] - _GRAD_ROW0)) < 1e-7
        assert abs(g.sum()) < 1e-12
    finally:
        minpy.set_global_policy('prefer_cuda')


def test_customop_rejects_unknown_variant():
    from minpy.primitive import customop
    with pytest.raises(ValueError):
        customop('tensorflow')


@pytest.mark.gpu
def test_customop_cuda_and_numpy_op_on_device_arrays():
    """(1) `customop('cuda')` ('mxnet' is accepted as its old name): the body receives RAW device arrays
    and composes library calls / CudaArray operators; (2) the reference's `customop('numpy')` softmax
    applied to DEVICE arrays under prefer_cuda: device->host hop in, gradient flows back through it."""
    import minpy
    minpy.set_global_policy('prefer_cuda')
    import minpy.numpy as mnp
    from minpy.core import grad_and_loss
    from minpy.primitive import customop
    from minpy_b200 import ops
    from minpy_b200.cuda_array import CudaArray

    @customop('mxnet')
    def softmax_dev(x, y):
        assert isinstance(x, CudaArray) and isinstance(y, CudaArray)
        probs = ops.softmax_rows(x)
        hot = ops.onehot_encode(y, ops.zeros(x.shape))
        return (ops.unary(ops.L().MC_LOG, probs) * hot).sum() * (-1.0 / x.shape[0])

    def softmax_dev_grad(ans, x, y):
        def grad(g):
            hot = ops.onehot_encode(y, ops.zeros(x.shape))
            return (ops.softmax_rows(x) - hot) * (float(g) / x.shape[0] if ops.is_scalar(g) else 1.0 / x.shape[0])
        return grad
    softmax_dev.def_grad(softmax_dev_grad)

    x, label = mnp.array(_X), mnp.array(_LABEL)
    assert x.on_device()
    for op in (softmax_dev, _softmax_numpy()):
        g, loss = grad_and_loss(op)(x, label)
        assert abs(float(loss) - _LOSS) < 1e-5 * _LOSS
        assert np.max(np.abs(g.asnumpy()[0] - _GRAD_ROW0)) < 1e-6
    # a device function BEHIND the host hop: d/dx sum-of-squares(2x) through the NumPy custom op
    softmax = _softmax_numpy()
    g2, _ = grad_and_loss(lambda v: softmax(v * 2.0, label))(x)
    ref = _softmax_numpy()
    minpy.set_global_policy('only_numpy')
    try:
        import minpy.numpy as hnp
        want, _ = grad_and_loss(lambda v: ref(v * 2.0, hnp.array(_LABEL)))(hnp.array(_X))
    finally:
        minpy.set_global_policy('prefer_cuda')
    assert np.max(np.abs(g2.asnumpy() - want.asnumpy())) < 1e-6


# ------------------------------------------------------------------------------- f3 -------
def _train(solver, batches, lo, hi):
    for i in range(lo, hi):
        solver._step(batches[i])  # pylint: disable=protected-access


@pytest.mark.parametrize('rule', ['sgd_momentum', 'adam'])
@pytest.mark.parametrize('graph', [False, True])
def test_checkpoint_resume_equals_uninterrupted_training(policy, rule, graph, tmp_path):
    """save_checkpoint after 3 steps, load into a FRESH solver, 3 more steps == 6 uninterrupted steps:
    parameters, batchnorm running statistics and optimizer state (velocity / m, v, t) all round-trip."""
    if graph and policy != 'only_cuda':
        pytest.skip('step graphs are a device feature')
    import minpy.numpy as mnp
    from minpy.nn.io import DataBatch
    from minpy.nn.solver import Solver
    from minpy_b200 import workloads
    rng = np.random.RandomState(5)
    batches = [DataBatch([mnp.array(rng.randn(32, 3, 4, 4).astype(np.float32))],
                         [mnp.array(rng.randint(0, 10, 32))], pad=0, index=i) for i in range(6)]
    cfg = {'learning_rate': 0.05, 'momentum': 0.9} if rule == 'sgd_momentum' else {'learning_rate': 1e-2}

    def make():
        model = workloads.DeepMLP(in_dim=48, hidden=(32, 32), classes=10, drop=0.0)
        solver = Solver(model, None, None, update_rule=rule, optim_config=dict(cfg), init_rule='gaussian',
                        init_config={'stdvar': 0.1}, verbose=False, step_graph=graph)
        solver.init()
        for k, v in workloads.init_params(model, 3, scale=0.1).items():
            model.params[k] = mnp.array(v)
        return model, solver

    model_a, solver_a = make()
    _train(solver_a, batches, 0, 6)
    model_b, solver_b = make()
    _train(solver_b, batches, 0, 3)
    solver_b.epoch = 2
    prefix = str(tmp_path / 'ckpt')
    solver_b.save_checkpoint(prefix)
    assert os.path.exists(prefix + '.params') and os.path.exists(prefix + '.solver')
    model_c, solver_c = make()
    solver_c.load_checkpoint(prefix)
    assert solver_c.epoch == 2 and len(solver_c.loss_history) == 3
    key = 'velocity' if rule == 'sgd_momentum' else 'm'
    for name in model_b.params:
        np.testing.assert_array_equal(np.asarray(solver_c.optim_configs[name][key].asnumpy(), np.float32),
                                      np.asarray(solver_b.optim_configs[name][key].asnumpy(), np.float32))
        if rule == 'adam':
            assert solver_c.optim_configs[name]['t'] == 3
    for name in model_b.aux_params:
        np.testing.assert_array_equal(np.asarray(model_c.aux_params[name].asnumpy(), np.float32),
                                      np.asarray(model_b.aux_params[name].asnumpy(), np.float32))
    _train(solver_c, batches, 3, 6)
    tol = 1e-6 if policy == 'only_numpy' else 2e-5      # float32 parameters; mixed-precision state on the host
    for name in model_a.params:
        a, c = model_a.params[name].asnumpy(), model_c.params[name].asnumpy()
        assert np.max(np.abs(a - c)) <= tol * max(1.0, np.max(np.abs(a))), name
    for name in model_a.aux_params:
        a, c = model_a.aux_params[name].asnumpy(), model_c.aux_params[name].asnumpy()
        assert np.max(np.abs(a - c)) <= tol * max(1.0, np.max(np.abs(a))), name
    if policy == 'only_cuda':
        assert all(v.on_device() for v in model_c.params.values())


def test_load_reports_a_missing_parameter(policy, tmp_path):
    import minpy.numpy as mnp
    from minpy.nn.model import ParamsNameNotFoundError
    from minpy_b200 import workloads
    small = workloads.MLP(in_dim=8, hidden=4, classes=3)
    small.params = dict((k, mnp.zeros(c['shape'])) for k, c in small.param_configs.items())
    del small.params['b2']
    small.save(str(tmp_path / 'm'))
    full = workloads.MLP(in_dim=8, hidden=4, classes=3)
    full.params = dict((k, mnp.zeros(c['shape'])) for k, c in full.param_configs.items())
    with pytest.raises(ParamsNameNotFoundError):
        full.load(str(tmp_path / 'm'))


# ------------------------------------------------------------------------------- f4 -------
def _stat_text(mnp):
    buf = io.StringIO()
    with redirect_stdout(buf):
        mnp.show_op_stat()
    return buf.getvalue()


def test_op_stat_counts_numpy_dispatch():
    import minpy
    minpy.set_global_policy('only_numpy')
    try:
        import minpy.numpy as mnp
        mnp.record_op_stat()
        a = mnp.ones((3, 4))
        b = mnp.cumsum(a * 2.0, axis=1)
        assert np.all(b.asnumpy() == np.cumsum(np.full((3, 4), 2.0), axis=1))
        text = _stat_text(mnp)
        assert text.startswith('--------Op Dispatch Statistics Start--------')
        assert 'NUMPY op called times:' in text and ' cumsum : 1' in text and ' multiply : 1' in text
        assert '100.0% in NumPy' in text.replace('numpy', 'NumPy') or '100.0%' in text
    finally:
        minpy.set_global_policy('prefer_cuda')


@pytest.mark.gpu
def test_prefer_cuda_falls_back_to_numpy_for_names_without_a_kernel():
    """The reference's promise (README.md:56-63, policy.py:220-231): under the preferring policy a name
    without a device implementation (where / cumsum / var / clip / sort ...) still works on a device
    array -- device->host copy, NumPy, the result wrapped and usable by device primitives again -- and
    the op statistics show which side served which call."""
    import minpy
    minpy.set_global_policy('prefer_cuda')
    import minpy.numpy as mnp
    rng = np.random.RandomState(9)
    h = rng.randn(16, 24).astype(np.float32)
    x = mnp.array(h) * 1.0
    assert x.on_device()
    np.testing.assert_allclose(mnp.where(x > 0.5, x, 0.0).asnumpy(), np.where(h > 0.5, h, 0.0), rtol=1e-6)
    np.testing.assert_allclose(mnp.cumsum(x, axis=1).asnumpy(), np.cumsum(h, axis=1), rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(mnp.var(x, axis=0).asnumpy(), np.var(h, axis=0), rtol=1e-5)
    np.testing.assert_allclose(mnp.clip(x, -0.3, 0.3).asnumpy(), np.clip(h, -0.3, 0.3), rtol=1e-6)
    np.testing.assert_allclose(mnp.sort(x, axis=1).asnumpy(), np.sort(h, axis=1), rtol=1e-6)
    # host result -> device primitive again (host->device copy on demand)
    c = mnp.cumsum(x, axis=1)
    s = mnp.sum(c * x)
    assert s.on_device()
    np.testing.assert_allclose(float(s), float(np.sum(np.cumsum(h, axis=1) * h)), rtol=1e-4)
    # op statistics, in the reference's format (docs/feature/op_stat.rst)
    mnp.record_op_stat()
    try:
        y = mnp.array(h) * 2.0
        mnp.cumsum(y, axis=0)
        mnp.sum(y)
        text = _stat_text(mnp)
        policy = minpy.get_global_policy()
        from minpy.array_variants import ArrayType
        assert policy._op_stat[ArrayType.NUMPY]['cumsum'] == 1          # pylint: disable=protected-access
        assert policy._op_stat[ArrayType.CUDA]['multiply'] == 1 and policy._op_stat[ArrayType.CUDA]['sum'] == 1  # pylint: disable=protected-access
        assert 'CUDA op called times:' in text and 'NUMPY op called times:' in text
        assert 'Total Dispatch Proportion:' in text
    finally:
        minpy.set_global_policy('prefer_cuda')
    # under only_cuda the same name is simply absent (no silent CPU execution on the product path)
    minpy.set_global_policy('only_cuda')
    try:
        import minpy.numpy as dnp
        assert not hasattr(dnp, 'cumsum')
    finally:
        minpy.set_global_policy('prefer_cuda')
