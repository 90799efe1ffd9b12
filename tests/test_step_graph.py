"""CUDA-graph replay of Solver._step (minpy_b200/step_graph.py) against the eager step: same
kernels, same inputs => the same parameters, optimizer state, running statistics and losses."""
import warnings

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.max(np.abs(a - b)) / max(1e-30, np.max(np.abs(b))))


def _run(make_model, rule, cfg, X, Y, batch, steps, graph, host_batches=False, lr_change_at=None, hook=None):
    import minpy
    minpy.set_global_policy('only_cuda')
    from minpy import array as marray
    from minpy.nn.io import DataBatch
    from minpy.nn.solver import Solver
    from minpy_b200 import workloads
    from minpy_b200.cuda_array import CudaArray
    import minpy.numpy.random as mrandom
    mrandom.seed(7)
    model = make_model()
    solver = Solver(model, None, None, update_rule=rule, optim_config=dict(cfg), init_rule='gaussian',
                    init_config={'stdvar': 0.01}, verbose=False, step_graph=graph)
    solver.init()
    for k, v in workloads.init_params(model, 3).items():
        model.params[k] = marray.wrap(CudaArray.from_numpy(v))
    nb = X.shape[0] // batch
    for i in range(steps):
        j = i % nb
        xb, yb = X[j * batch:(j + 1) * batch], Y[j * batch:(j + 1) * batch]
        if not host_batches:
            xb, yb = marray.wrap(CudaArray.from_numpy(xb)), marray.wrap(CudaArray.from_numpy(yb))
        if lr_change_at is not None and i == lr_change_at:
            for c in solver.optim_configs.values():
                c['learning_rate'] *= 0.5
        if hook:
            hook(solver, i)
        solver._step(DataBatch([xb], [yb], pad=0, index=i))  # pylint: disable=protected-access
    params = dict((k, v.asnumpy()) for k, v in model.params.items())
    aux = dict((k, v.asnumpy()) for k, v in model.aux_params.items() if hasattr(v, 'asnumpy'))
    state = {}
    for name, c in solver.optim_configs.items():
        for k, v in c.items():
            state[(name, k)] = v.asnumpy() if hasattr(v, 'asnumpy') else v
    losses = [float(l) for l in solver.loss_history]
    return params, aux, state, losses, solver


def _same(a, b, tol):
    pa, xa, sa, la, _ = a
    pb, xb, sb, lb, _ = b
    assert len(la) == len(lb)
    assert rel(la, lb) <= tol, (la, lb)
    for k in pb:
        assert rel(pa[k], pb[k]) <= tol, k
    for k in xb:
        assert rel(xa[k], xb[k]) <= tol, k
    assert set(sa) == set(sb)
    for k in sb:
        assert rel(sa[k], sb[k]) <= tol, k


def _mlp():
    from minpy_b200 import workloads
    return workloads.MLP(in_dim=96, hidden=64, classes=10)


def _bn_mlp():
    from minpy_b200 import workloads
    return workloads.DeepMLP(in_dim=96, hidden=(64, 64), classes=10, drop=0.0)


def _data(n, d=96, seed=0):
    rng = np.random.RandomState(seed)
    return rng.randn(n, d).astype(np.float32), rng.randint(0, 10, size=n)


@pytest.mark.parametrize('rule,cfg', [('sgd', {'learning_rate': 0.1}),
                                      ('sgd_momentum', {'learning_rate': 0.05, 'momentum': 0.9}),
                                      ('rmsprop', {'learning_rate': 1e-3}),
                                      ('adam', {'learning_rate': 1e-3})])
def test_graph_replay_equals_eager(rule, cfg):
    """8 steps on 4 distinct batches: replay reads the new batch, the updated parameters and the
    updated optimizer state (adam: the advancing t) -- bit-for-bit the eager kernels."""
    X, Y = _data(4 * 32)
    eager = _run(_mlp, rule, cfg, X, Y, 32, 8, False)
    graph = _run(_mlp, rule, cfg, X, Y, 32, 8, True)
    assert graph[4]._step_graph.replays == 6 and graph[4]._step_graph.captures == 1  # pylint: disable=protected-access
    _same(graph, eager, 1e-6)
    if rule == 'adam':
        assert graph[2][('w1', 't')] == eager[2][('w1', 't')] == 8


def test_graph_replay_batchnorm_running_stats_and_host_batches():
    """BatchNorm auxiliary state is rebound by the model every step; host (NumPy) batches are
    copied into the static input buffers."""
    X, Y = _data(3 * 64, seed=1)
    cfg = {'learning_rate': 0.05, 'momentum': 0.9}
    eager = _run(_bn_mlp, 'sgd_momentum', cfg, X, Y, 64, 7, False)
    graph = _run(_bn_mlp, 'sgd_momentum', cfg, X.astype(np.float64), Y, 64, 7, True, host_batches=True)
    assert graph[4]._step_graph.replays == 5   # pylint: disable=protected-access
    assert len(eager[1]) == 4
    _same(graph, eager, 1e-6)


def test_graph_recaptures_when_learning_rate_changes():
    X, Y = _data(2 * 32, seed=2)
    cfg = {'learning_rate': 0.05, 'momentum': 0.9}
    eager = _run(_mlp, 'sgd_momentum', cfg, X, Y, 32, 8, False, lr_change_at=5)
    graph = _run(_mlp, 'sgd_momentum', cfg, X, Y, 32, 8, True, lr_change_at=5)
    assert graph[4]._step_graph.captures == 2   # pylint: disable=protected-access
    _same(graph, eager, 1e-6)


def test_graph_dropout_masks_differ_between_replays():
    """Same batch every step, learning rate 0: the only thing that changes the loss from replay to
    replay is the dropout mask, which must be a new draw each time."""
    from minpy_b200 import workloads
    X, Y = _data(64, seed=3)
    make = lambda: workloads.DeepMLP(in_dim=96, hidden=(64, 64), classes=10, drop=0.5)
    out = _run(make, 'sgd', {'learning_rate': 0.0}, X, Y, 64, 8, True)
    losses = out[3]
    assert out[4]._step_graph.replays == 6   # pylint: disable=protected-access
    assert all(np.isfinite(losses))
    assert len(set(round(l, 7) for l in losses[2:])) == 6, losses


def test_graph_falls_back_when_the_step_reads_the_device():
    """A loss function that fetches a device value cannot be captured: the solver warns once and
    keeps training eagerly with correct results."""
    from minpy_b200 import workloads

    class Peeking(workloads.MLP):
        def loss(self, predict, y):
            l = super(Peeking, self).loss(predict, y)
            float(l)                     # device -> host read inside the step
            return l

    make = lambda: Peeking(in_dim=96, hidden=64, classes=10)
    X, Y = _data(2 * 32, seed=4)
    cfg = {'learning_rate': 0.05, 'momentum': 0.9}
    eager = _run(make, 'sgd_momentum', cfg, X, Y, 32, 6, False)
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter('always')
        graph = _run(make, 'sgd_momentum', cfg, X, Y, 32, 6, True)
    assert graph[4]._step_graph.disabled   # pylint: disable=protected-access
    assert any('step graph disabled' in str(x.message) for x in w)
    _same(graph, eager, 1e-6)


def test_graph_replay_conv_trunk():
    """The symbol path (core.Function conv trunk) replays too."""
    from minpy_b200 import workloads
    rng = np.random.RandomState(5)
    X = rng.randn(3 * 16, 3, 32, 32).astype(np.float32)
    Y = rng.randint(0, 10, size=3 * 16)
    make = lambda: workloads.ConvNet(batch=16, filters=(8, 16, 32))
    cfg = {'learning_rate': 0.01, 'momentum': 0.9}
    eager = _run(make, 'sgd_momentum', cfg, X, Y, 16, 6, False)
    graph = _run(make, 'sgd_momentum', cfg, X, Y, 16, 6, True)
    assert graph[4]._step_graph.replays == 4   # pylint: disable=protected-access
    _same(graph, eager, 1e-6)


def test_graph_replay_lstm_adam():
    from minpy_b200 import workloads
    make = lambda: workloads.CharLSTM(vocab=16, hidden=32)
    X, Y = workloads.char_like(2 * 8, T=6, V=16)
    eager = _run(make, 'adam', {'learning_rate': 1e-2}, X, Y, 8, 6, False)
    graph = _run(make, 'adam', {'learning_rate': 1e-2}, X, Y, 8, 6, True)
    assert graph[4]._step_graph.replays == 4   # pylint: disable=protected-access
    _same(graph, eager, 1e-5)


def test_graph_falls_back_when_the_step_uploads_a_host_constant():
    """A loss that builds a device array from a NumPy constant inside the step: the host->device copy
    must not become a graph node (every replay would re-read a freed host buffer): the capture is
    refused (mc_memcpy_h2d requires !capturing) and training continues eagerly, correctly."""
    import minpy.numpy as mnp
    from minpy_b200 import workloads
    weights = np.linspace(0.5, 1.5, 10).astype(np.float32)

    class Weighted(workloads.MLP):
        def loss(self, predict, y):
            return super(Weighted, self).loss(predict * mnp.array(weights), y)

    make = lambda: Weighted(in_dim=96, hidden=64, classes=10)
    X, Y = _data(2 * 32, seed=6)
    cfg = {'learning_rate': 0.05, 'momentum': 0.9}
    eager = _run(make, 'sgd_momentum', cfg, X, Y, 32, 6, False)
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter('always')
        graph = _run(make, 'sgd_momentum', cfg, X, Y, 32, 6, True)
    assert graph[4]._step_graph.disabled   # pylint: disable=protected-access
    assert any('step graph disabled' in str(x.message) for x in w)
    _same(graph, eager, 1e-6)


def test_graph_adam_outside_the_multi_tensor_kernel_is_refused():
    """A parameter the loss does not reach has gradient 0.0: optim.multi_update declines and the
    per-parameter mc_adam (bias correction baked from the host-side t) would be captured -- the step
    graph must refuse rather than replay a stale t."""
    from minpy_b200 import workloads

    class Unused(workloads.MLP):
        def __init__(self, **kw):
            super(Unused, self).__init__(**kw)
            self.add_param(name='unused', shape=(4,))

    make = lambda: Unused(in_dim=96, hidden=64, classes=10)
    X, Y = _data(2 * 32, seed=7)
    eager = _run(make, 'adam', {'learning_rate': 1e-2}, X, Y, 32, 6, False)
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter('always')
        graph = _run(make, 'adam', {'learning_rate': 1e-2}, X, Y, 32, 6, True)
    assert graph[4]._step_graph.disabled   # pylint: disable=protected-access
    assert any('multi-tensor' in str(x.message) for x in w)
    _same(graph, eager, 1e-6)


def test_graph_training_keeps_best_params_and_fresh_host_views():
    """Solver.train with step_graph=True: replays write parameters IN PLACE, so (a)
    `best_params` must be a snapshot, not an alias of the live buffers, and (b) a host copy cached by
    an earlier asnumpy() must not be served after later replays."""
    import minpy
    minpy.set_global_policy('only_cuda')
    import minpy.numpy as mnp
    from minpy.nn.io import NDArrayIter
    from minpy.nn.solver import Solver
    from minpy_b200 import workloads
    rng = np.random.RandomState(8)
    X = rng.randn(256, 96).astype(np.float32)
    W = rng.randn(96, 10).astype(np.float32)
    Y = np.argmax(X.dot(W), axis=1)
    train = NDArrayIter(X, Y, batch_size=32, shuffle=False)
    # the validation set is RANDOM labels scored after a huge-lr epoch: accuracy moves around, so the
    # best epoch is (almost surely) not the last one
    test = NDArrayIter(X[:64], rng.randint(0, 10, 64), batch_size=32, shuffle=False)
    model = workloads.MLP(in_dim=96, hidden=64, classes=10)
    solver = Solver(model, train, test, update_rule='sgd_momentum', optim_config={'learning_rate': 0.5, 'momentum': 0.9},
                    init_rule='gaussian', init_config={'stdvar': 0.05}, num_epochs=6, verbose=False, step_graph=True)
    solver.init()
    seen = []
    orig = solver.check_accuracy

    def spy(it, num_samples=None):
        seen.append(dict((k, v.asnumpy().copy()) for k, v in model.params.items()))   # caches a host copy
        return orig(it, num_samples=num_samples)
    solver.check_accuracy = spy
    solver.train()
    sg = solver._step_graph  # pylint: disable=protected-access
    assert sg is not None and not sg.disabled and sg.replays > 20
    # (b) the per-epoch host views differ from epoch to epoch (they would all equal the first one if
    # the cached NumPy copy were served)
    per_epoch = seen[0::2]
    assert len(per_epoch) == 6
    for a, b in zip(per_epoch[:-1], per_epoch[1:]):
        assert np.max(np.abs(a['w1'] - b['w1'])) > 0
    # (a) best_params equals the snapshot of the best epoch, not necessarily the last
    best_epoch = int(np.argmax(solver.val_acc_history))       # first maximum: `>` keeps the earliest
    for k in model.params:
        assert np.array_equal(model.params[k].asnumpy(), per_epoch[best_epoch][k]), k
    if best_epoch != 5:
        assert np.max(np.abs(model.params['w1'].asnumpy() - per_epoch[5]['w1'])) > 0


def test_what_a_script_observes_under_the_step_graph():
    """what reference-style script code sees when `step_graph=True`.
    (a) `solver.loss_history[-1]` read right after a step is THAT step's loss, as solver.py:187 gives it
        (the entry is a lazy device scalar; float() waits for its own copy only);
    (b) parameters: the reference REBINDS model.params[k] every step (solver.py:190-195), so an array a
        script held from before the step keeps the old values.  Eager stepping here does the same.  A
        replayed step writes the parameter buffers IN PLACE: the held array IS the live
        parameter and shows the new values.  Documented difference of the opt-in mode; this test pins it."""
    from minpy_b200 import workloads
    make = lambda: workloads.MLP(in_dim=96, hidden=64, classes=10)
    X, Y = _data(4 * 32, seed=11)
    seen = {False: [], True: []}
    held = {}

    def hook_for(graph):
        def hook(solver, i):
            if i > 0:
                seen[graph].append(float(solver.loss_history[-1]))        # the step just taken
            if i == 4:
                held[graph] = (solver.model.params['w1'], solver.model.params['w1'].asnumpy().copy())
        return hook
    eager = _run(make, 'sgd_momentum', {'learning_rate': 0.05, 'momentum': 0.9}, X, Y, 32, 6, False, hook=hook_for(False))
    graph = _run(make, 'sgd_momentum', {'learning_rate': 0.05, 'momentum': 0.9}, X, Y, 32, 6, True, hook=hook_for(True))
    assert not graph[4]._step_graph.disabled and graph[4]._step_graph.replays >= 3   # pylint: disable=protected-access
    assert rel(seen[True], seen[False]) <= 1e-6 and rel(seen[False], eager[3][:-1]) == 0.0
    _same(graph, eager, 1e-6)
    obj, before = held[False]
    assert np.array_equal(obj.asnumpy(), before)                 # eager: the held array kept the old values
    assert eager[4].model.params['w1'] is not obj
    obj, before = held[True]
    assert graph[4].model.params['w1'] is obj                     # graph: the held array is the live parameter
    assert rel(obj.asnumpy(), graph[0]['w1']) == 0.0 and not np.array_equal(obj.asnumpy(), before)
