"""x1 (north_star): the reference's OWN example scripts -- examples/nn/mlp.py, mlp_bn_dropout.py,
cnn.py, lstm.py, byte-for-byte as oracle/build_ref.py copied them into oracle/_ref/examples/nn
(or straight from /root/reference when that is present) -- run UNMODIFIED against this repository's
`minpy` (+ the `mxnet` symbol facade for cnn.py).

Only the data source is replaced, as SURVEY 8d prescribes (no CIFAR files, no network):
`examples.utils.data_utils` is a stand-in module whose `get_CIFAR10_data` / `adding_problem_generator`
return small synthetic sets of the same shapes and dtypes (the original imports `scipy.misc.imread`,
which modern SciPy no longer has).  `xrange` (lstm.py:29, Python 2) is aliased to `range`.

* CPU (`only_numpy` policy): mlp.py, mlp_bn_dropout.py, lstm.py train for their 10 epochs on the NumPy
  variant of the re-created front end.
* GPU (default `prefer_cuda` policy, nothing set by the test): all four scripts train on the cuda
  variant; the library's launch counter must show the steps ran as CUDA kernels and the parameters
  must live on the device.
Each run must produce a finite, decreasing loss and record validation accuracies.
"""
import builtins
import os
import runpy
import sys
import types

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _script_dir():
    for d in (os.path.join(ROOT, 'oracle', '_ref', 'examples', 'nn'), '/root/reference/examples/nn'):
        if os.path.isfile(os.path.join(d, 'mlp.py')):
            return d
    return None


needs_scripts = pytest.mark.skipif(_script_dir() is None, reason='reference example scripts not present '
                                   '(oracle/_ref is assembled by oracle/build_ref.py in the build container)')


def _data_utils():
    """Stand-in for examples/utils/data_utils.py:37-79,191-230: same keys, shapes and dtypes."""
    mod = types.ModuleType('examples.utils.data_utils')

    def get_CIFAR10_data(data_dir, num_training=512, num_validation=128, num_test=128):  # pylint: disable=unused-argument
        rng = np.random.RandomState(0)
        proj = rng.randn(3 * 32 * 32, 10)

        def make(n):
            x = rng.randn(n, 3, 32, 32)
            return x, np.argmax(x.reshape(n, -1).dot(proj), axis=1)          # learnable labels
        xt, yt = make(num_training)
        xv, yv = make(num_validation)
        xs, ys = make(num_test)
        return {'X_train': xt, 'y_train': yt, 'X_val': xv, 'y_val': yv, 'X_test': xs, 'y_test': ys}

    def adding_problem_generator(N, seq_len=30, high=1):  # pylint: disable=unused-argument
        n = min(int(N), 300)
        rng = np.random.RandomState(n)
        x_num = rng.uniform(0, high, size=(n, seq_len, 1))
        x_mask = np.zeros((n, seq_len, 1))
        y = np.ones((n, 1))
        for i in range(n):
            pos = rng.choice(seq_len, size=2, replace=False)
            x_mask[i, pos, 0] = 1
            y[i, 0] = np.sum(x_num[i, pos, 0])
        return np.append(x_num, x_mask, axis=2), y

    mod.get_CIFAR10_data = get_CIFAR10_data
    mod.adding_problem_generator = adding_problem_generator
    return mod


def _run_script(name, policy):
    """runpy the script as __main__; returns the Solver it trained."""
    import minpy
    from minpy.nn import solver as solver_mod
    if policy is not None:
        minpy.set_global_policy(policy)
    pkgs = {'examples': types.ModuleType('examples'), 'examples.utils': types.ModuleType('examples.utils'),
            'examples.utils.data_utils': _data_utils()}
    pkgs['examples'].utils = pkgs['examples.utils']
    pkgs['examples.utils'].data_utils = pkgs['examples.utils.data_utils']
    saved_mods = dict((k, sys.modules.get(k)) for k in pkgs)
    saved_argv, had_xrange = sys.argv, hasattr(builtins, 'xrange')
    trained = []
    orig_train = solver_mod.Solver.train

    def train(self, *a, **k):
        trained.append(self)
        return orig_train(self, *a, **k)
    solver_mod.Solver.train = train
    sys.modules.update(pkgs)
    sys.argv = [name, '--data_dir', 'synthetic']
    if not had_xrange:
        builtins.xrange = range
    np.random.seed(0)
    try:
        runpy.run_path(os.path.join(_script_dir(), name), run_name='__main__')
    finally:
        solver_mod.Solver.train = orig_train
        sys.argv = saved_argv
        for k, v in saved_mods.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
        if not had_xrange:
            del builtins.xrange
    assert len(trained) == 1
    return trained[0]


def _check_trained(solver, epochs=10):
    losses = np.array([float(np.asarray(l.asnumpy() if hasattr(l, 'asnumpy') else l)) for l in solver.loss_history])
    assert losses.size >= epochs and np.all(np.isfinite(losses))
    k = max(2, losses.size // 5)
    assert losses[-k:].mean() < losses[:k].mean(), (losses[:k], losses[-k:])
    assert solver.epoch == epochs
    assert len(solver.val_acc_history) == epochs and len(solver.train_acc_history) == epochs


@needs_scripts
@pytest.mark.parametrize('name', ['mlp.py', 'mlp_bn_dropout.py', 'lstm.py'])
def test_reference_script_runs_on_the_numpy_variant(name):
    import minpy
    try:
        solver = _run_script(name, 'only_numpy')
        _check_trained(solver)
        from minpy.array_variants import ArrayType
        for value in solver.model.params.values():
            assert value.has_type(ArrayType.NUMPY) and not value.has_type(ArrayType.CUDA)
    finally:
        minpy.set_global_policy('prefer_cuda')


@needs_scripts
@pytest.mark.gpu
@pytest.mark.parametrize('name', ['mlp.py', 'mlp_bn_dropout.py', 'cnn.py', 'lstm.py'])
def test_reference_script_runs_on_the_cuda_variant(name):
    """Default policy, as a user of the reference would find it: nothing but `python script.py`."""
    import ctypes
    import minpy
    from minpy_b200._lib import get_lib
    minpy.set_global_policy('prefer_cuda')
    lib = get_lib()
    n0, n1 = ctypes.c_uint64(), ctypes.c_uint64()
    lib.mc_launch_count(ctypes.byref(n0))
    solver = _run_script(name, None)
    lib.mc_launch_count(ctypes.byref(n1))
    _check_trained(solver)
    steps = len(solver.loss_history)
    assert n1.value - n0.value >= 10 * steps, (n1.value - n0.value, steps)     # the steps ran as CUDA kernels
    for value in solver.model.params.values():
        assert value.on_device()
