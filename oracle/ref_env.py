"""ORACLE tooling -- how the UNMODIFIED reference is imported in a modern interpreter.

Test / measurement infrastructure only: used by oracle/gen_golden.py (golden vectors, straight from
/root/reference), oracle/build_ref.py (assembles oracle/_ref/, the copy that travels to the GPU box),
oracle/ref_runner.py (the timed CPU baseline) and tests/.  Nothing on the product path imports it.

What `activate(prefix)` does (the reference's sources are never edited):
  * a stub `mxnet` package (version string, Context, ndarray.NDArray, the op names
    array_variants/mxnet/mxnet_core.py:112-170 looks up at import time) and an empty `h5py` are put
    FIRST on sys.path -- they only satisfy `import`, every call raises;
  * aliases the 2017 sources use and modern Python / NumPy removed are restored in this process only:
    numpy.float/int/bool (array_variants/numpy/__init__.py:11-14), collections.Iterable
    (array.py:337), builtins.xrange (numpy_core.py:107, examples/nn/lstm.py:29);
  * `prefix` (the directory that CONTAINS the reference's `minpy` package) goes on sys.path and the
    repository root -- which holds the new `minpy` -- is taken off it (or moved behind, `keep_root`);
  * the policy is set to 'only_numpy', so every number comes from the reference's own NumPy variant,
    tape, layers, optim and solver code.
"""
from __future__ import print_function

import builtins
import collections
import collections.abc
import os
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF_BUILT = os.path.join(HERE, '_ref')            # written by oracle/build_ref.py (git-ignored)

MXNET_STUB = '''
__version__ = "0.9.2"
import types, sys
class Context(object):
    def __init__(self, device_type="cpu", device_id=0):
        self.device_type, self.device_id = device_type, device_id
    def __enter__(self): return self
    def __exit__(self, *a): return False
def cpu(i=0): return Context("cpu", i)
def gpu(i=0): return Context("gpu", i)
def _op(*a, **k): raise RuntimeError("mxnet stub: no arithmetic here")
class _Mod(types.ModuleType):
    def __getattr__(self, name):
        if name.startswith("__"): raise AttributeError(name)
        return _op
class NDArray(object): pass
ndarray = _Mod("mxnet.ndarray"); ndarray.NDArray = NDArray
nd = ndarray
random = _Mod("mxnet.random")
symbol = _Mod("mxnet.symbol"); sym = symbol
class _Sym(object): pass
symbol.Symbol = _Sym
io = _Mod("mxnet.io")
class DataIter(object): pass
class DataBatch(object): pass
io.DataIter = DataIter; io.DataBatch = DataBatch
def _mk(name):
    def f(*a, **k): raise RuntimeError("mxnet stub: %s has no arithmetic" % name)
    f.__name__ = name
    return f
for _name in ("abs add ceil cos divide dot exp expand_dims floor log max maximum min minimum multiply negative "
              "power round sign sin sqrt subtract sum tanh transpose true_divide").split():
    ndarray.__dict__[_name] = _mk(_name)
for _name in ("normal", "uniform"):
    random.__dict__[_name] = _mk(_name)
NDArray.reshape = _mk("reshape")
for _n, _m in (("ndarray", ndarray), ("nd", nd), ("random", random), ("symbol", symbol), ("io", io)):
    sys.modules["mxnet." + _n] = _m
'''


def write_stubs(dest):
    """Write the stub `mxnet` and `h5py` packages under `dest`; returns dest."""
    for pkg, text in (('mxnet', MXNET_STUB), ('h5py', '')):
        os.makedirs(os.path.join(dest, pkg), exist_ok=True)
        with open(os.path.join(dest, pkg, '__init__.py'), 'w') as f:
            f.write(text)
    return dest


def apply_shims():
    import numpy as onp
    for alias, target in (('float', float), ('int', int), ('bool', bool), ('object', object), ('complex', complex)):
        if not hasattr(onp, alias):
            setattr(onp, alias, target)
    for name in ('Iterable', 'Mapping', 'Sequence', 'Callable'):
        if not hasattr(collections, name):
            setattr(collections, name, getattr(collections.abc, name))
    builtins.xrange = range


def built_prefix():
    """oracle/_ref/site when oracle/build_ref.py has run, else None."""
    site = os.path.join(REF_BUILT, 'site')
    return site if os.path.isdir(os.path.join(site, 'minpy')) else None


def activate(prefix=None, keep_root=False, stubs=None):
    """Import the reference's `minpy` from `prefix` under the 'only_numpy' policy; returns the module."""
    if prefix is None:
        prefix = built_prefix()
        if prefix is None:
            raise RuntimeError('oracle/_ref is not built (python oracle/build_ref.py, build container only)')
    if stubs is None:
        stubs = os.path.join(REF_BUILT, 'stubs')
        if not os.path.isdir(os.path.join(stubs, 'mxnet')):
            stubs = write_stubs(tempfile.mkdtemp(prefix='minpy_ref_stubs_'))
    apply_shims()
    if 'minpy' in sys.modules:
        raise RuntimeError('a `minpy` package is already imported in this process; the reference needs its own')
    sys.path[:] = [p for p in sys.path if os.path.abspath(p or '.') not in (ROOT, HERE)]
    sys.path.insert(0, prefix)
    sys.path.insert(0, stubs)
    if keep_root:
        sys.path.append(ROOT)       # behind the reference: `minpy` and `mxnet` resolve to the reference / the stub
    import warnings
    warnings.simplefilter('ignore')
    import minpy
    assert os.path.abspath(minpy.__file__).startswith(os.path.abspath(prefix)), minpy.__file__
    minpy.set_global_policy('only_numpy')
    return minpy
