"""f2 remainder: the CIFAR-10 loader the example scripts import
(`examples.utils.data_utils.get_CIFAR10_data`, reference examples/utils/data_utils.py:7-79) on files
written in the published CIFAR-10 pickle format, and NDArrayIter's construction-time shuffle
(nn/io.py:172-177) on DEVICE-resident arrays (a row gather on the GPU, no host round trip)."""
import os
import pickle
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _write_cifar(root, per_batch=40, seed=0):
    rng = np.random.RandomState(seed)
    raw = {}
    for name in ['data_batch_%d' % b for b in range(1, 6)] + ['test_batch']:
        data = rng.randint(0, 256, size=(per_batch, 3072)).astype(np.uint8)
        labels = [int(v) for v in rng.randint(0, 10, per_batch)]
        with open(os.path.join(root, name), 'wb') as f:
            pickle.dump({'data': data, 'labels': labels, 'batch_label': name, 'filenames': []}, f, protocol=2)
        raw[name] = (data, np.array(labels))
    return raw


def test_cifar10_loader_on_the_published_format(tmp_path):
    from examples.utils.data_utils import get_CIFAR10_data, load_CIFAR10
    raw = _write_cifar(str(tmp_path))
    Xtr, ytr, Xte, yte = load_CIFAR10(str(tmp_path))
    assert Xtr.shape == (200, 32, 32, 3) and Xtr.dtype == np.float64 and Xte.shape == (40, 32, 32, 3)
    # row layout of the format: 1024 red, 1024 green, 1024 blue, each row-major 32x32
    d0, l0 = raw['data_batch_1']
    assert Xtr[3, 5, 7, 1] == d0[3, 1024 + 5 * 32 + 7] and ytr[3] == l0[3]
    assert Xtr[40 + 2, 31, 0, 2] == raw['data_batch_2'][0][2, 2048 + 31 * 32] and yte[9] == raw['test_batch'][1][9]
    data = get_CIFAR10_data(str(tmp_path), num_training=150, num_validation=30, num_test=20)
    assert data['X_train'].shape == (150, 3, 32, 32) and data['X_val'].shape == (30, 3, 32, 32)
    assert data['X_test'].shape == (20, 3, 32, 32) and data['y_val'].shape == (30,)
    mean = Xtr[:150].mean(axis=0)
    np.testing.assert_allclose(data['X_train'], (Xtr[:150] - mean).transpose(0, 3, 1, 2))
    np.testing.assert_allclose(data['X_val'], (Xtr[150:180] - mean).transpose(0, 3, 1, 2))
    np.testing.assert_allclose(data['X_test'], (Xte[:20] - mean).transpose(0, 3, 1, 2))
    assert np.abs(data['X_train'].mean(axis=0)).max() < 1e-9
    assert np.array_equal(data['y_train'], ytr[:150]) and data['X_train'].flags['C_CONTIGUOUS']


def test_synthetic_generators():
    from examples.utils.data_utils import adding_problem_generator, gaussian_cluster_generator
    np.random.seed(0)
    X, Y = adding_problem_generator(50, seq_len=12)
    assert X.shape == (50, 12, 2) and Y.shape == (50, 1)
    assert np.all(X[:, :, 1].sum(axis=1) == 2)
    np.testing.assert_allclose((X[:, :, 0] * X[:, :, 1]).sum(axis=1), Y[:, 0])
    x, y = gaussian_cluster_generator(100, 8, 5)
    assert x.shape == (100, 8) and y.shape == (100, 5) and np.all(y.sum(axis=1) == 1)


def test_mlp_script_trains_on_loader_output(tmp_path):
    """The loader's dictionary feeds NDArrayIter + Solver exactly as examples/nn/mlp.py:49-78 does."""
    import minpy
    minpy.set_global_policy('only_numpy')
    try:
        from examples.utils.data_utils import get_CIFAR10_data
        from minpy.nn.io import NDArrayIter
        from minpy.nn.solver import Solver
        from minpy_b200 import workloads
        _write_cifar(str(tmp_path), per_batch=32, seed=1)
        data = get_CIFAR10_data(str(tmp_path), num_training=128, num_validation=32, num_test=32)
        train = NDArrayIter(data['X_train'], data['y_train'], batch_size=32, shuffle=True)
        test = NDArrayIter(data['X_test'], data['y_test'], batch_size=32, shuffle=False)
        solver = Solver(workloads.MLP(hidden=16), train, test, num_epochs=2, init_rule='gaussian',
                        init_config={'stdvar': 0.001}, update_rule='sgd_momentum',
                        optim_config={'learning_rate': 1e-6, 'momentum': 0.9}, verbose=False)
        solver.init()
        solver.train()
        assert len(solver.loss_history) == 8 and np.all(np.isfinite([float(l) for l in solver.loss_history]))
    finally:
        minpy.set_global_policy('prefer_cuda')


@pytest.mark.gpu
def test_shuffle_of_device_resident_data_stays_on_the_device():
    import ctypes
    import minpy
    minpy.set_global_policy('only_cuda')
    try:
        import minpy.numpy as mnp
        from minpy.nn.io import NDArrayIter
        from minpy_b200._lib import get_lib
        rng = np.random.RandomState(3)
        X = rng.randn(96, 3, 4, 4).astype(np.float32)
        Y = np.arange(96)
        X[:, 0, 0, 0] = Y                                   # each row carries its label
        dx, dy = mnp.array(X), mnp.array(Y)
        assert dx.on_device() and dy.on_device()
        np.random.seed(5)
        it = NDArrayIter(dx, dy, batch_size=32, shuffle=True, num_parts=1, part_index=0)
        seen = []
        for b in it:
            assert b.data[0].on_device() and b.label[0].on_device()
            xb, yb = b.data[0].asnumpy(), b.label[0].asnumpy()
            assert np.array_equal(xb[:, 0, 0, 0].astype(np.int64), yb.astype(np.int64))   # rows and labels moved together
            assert np.array_equal(xb, X[yb.astype(np.int64)])
            seen.extend(int(v) for v in yb)
        assert sorted(seen) == list(range(96)) and seen != list(range(96))
        n = ctypes.c_uint64()
        get_lib().mc_launch_count(ctypes.byref(n))
        assert n.value > 0
    finally:
        minpy.set_global_policy('prefer_cuda')
