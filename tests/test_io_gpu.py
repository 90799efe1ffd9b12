"""NDArrayIter(device_prefetch=True): the batches copied ahead on the second stream are the batches
the plain iterator yields (nn/io.py:136-289 semantics: order, wrap-around padding, epochs)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_prefetched_batches_equal_host_batches():
    import minpy
    minpy.set_global_policy('only_cuda')
    from minpy.nn.io import NDArrayIter
    rng = np.random.RandomState(0)
    X = rng.randn(70, 3, 4).astype(np.float64)           # float64 host data -> float32 on the device
    Y = rng.randint(0, 10, size=70)
    plain = NDArrayIter(X, Y, batch_size=16, shuffle=False, num_parts=1, part_index=0)
    ahead = NDArrayIter(X, Y, batch_size=16, shuffle=False, num_parts=1, part_index=0, device_prefetch=True)
    for _epoch in range(3):
        plain.reset()
        ahead.reset()
        a_batches = []
        for b in ahead:
            # fetch immediately: a slot is recycled two batches later
            a_batches.append((b.data[0].asnumpy().copy(), b.label[0].asnumpy().copy(), b.pad))
        p_batches = [(b.data[0].asnumpy(), b.label[0].asnumpy(), b.pad) for b in plain]
        assert len(a_batches) == len(p_batches) == 5
        for (ax, ay, apad), (px, py, ppad) in zip(a_batches, p_batches):
            assert apad == ppad
            assert ax.dtype == np.float32 and ax.shape == px.shape
            assert np.array_equal(ax, px.astype(np.float32))
            assert np.array_equal(ay, py)
        # the next epoch's first batch is already on its way when the epoch ends
        assert ahead._prefetch.cursor == 0 and ahead._prefetch.pending is not None


def test_prefetch_roll_over_does_not_speculate():
    """'roll_over' starts the next epoch at a data-dependent cursor: no batch is sent ahead across
    the epoch boundary, and the batches still equal the plain iterator's."""
    import minpy
    minpy.set_global_policy('only_cuda')
    from minpy.nn.io import NDArrayIter
    rng = np.random.RandomState(1)
    X = rng.randn(40, 5).astype(np.float32)
    Y = rng.randint(0, 3, size=40)
    kw = dict(batch_size=16, shuffle=False, num_parts=1, part_index=0, last_batch_handle='roll_over')
    plain, ahead = NDArrayIter(X, Y, **kw), NDArrayIter(X, Y, device_prefetch=True, **kw)
    for _epoch in range(3):
        plain.reset()
        ahead.reset()
        a = [(b.data[0].asnumpy().copy(), b.label[0].asnumpy().copy()) for b in ahead]
        p = [(b.data[0].asnumpy(), b.label[0].asnumpy()) for b in plain]
        assert len(a) == len(p)
        for (ax, ay), (px, py) in zip(a, p):
            assert np.array_equal(ax, px) and np.array_equal(ay, py)


def test_prefetch_feeds_the_step_graph():
    """Solver.train() over a prefetching iterator with the step graph on: same parameters as the plain
    iterator with eager steps."""
    import minpy
    minpy.set_global_policy('only_cuda')
    from minpy import array as marray
    from minpy.nn.io import NDArrayIter
    from minpy.nn.solver import Solver
    from minpy_b200 import workloads
    from minpy_b200.cuda_array import CudaArray
    rng = np.random.RandomState(1)
    X = rng.randn(128, 96).astype(np.float32)
    Y = rng.randint(0, 10, size=128)

    def run(prefetch, graph):
        it = NDArrayIter(X, Y, batch_size=32, shuffle=False, num_parts=1, part_index=0, device_prefetch=prefetch)
        model = workloads.MLP(in_dim=96, hidden=64, classes=10)
        solver = Solver(model, it, it, update_rule='sgd_momentum', optim_config={'learning_rate': 0.05, 'momentum': 0.9},
                        init_rule='gaussian', init_config={'stdvar': 0.01}, verbose=False, num_epochs=2, step_graph=graph)
        solver.init()
        for k, v in workloads.init_params(model, 3).items():
            model.params[k] = marray.wrap(CudaArray.from_numpy(v))
        solver.train()
        return dict((k, v.asnumpy()) for k, v in model.params.items()), [float(l) for l in solver.loss_history]

    pa, la = run(False, False)
    pb, lb = run(True, True)
    assert len(la) == len(lb) == 8
    assert np.allclose(la, lb, rtol=1e-5, atol=1e-7)
    for k in pa:
        assert np.allclose(pa[k], pb[k], rtol=1e-5, atol=1e-7), k
