"""`mxnet` name façade for unmodified MinPy scripts.

The reference's CNN examples do `import mxnet as mx` and build `mx.sym.*` graphs that
`minpy.core.Function` runs (examples/nn/cnn.py:7,32-44).  MXNet is not a dependency of this
back end: this package only re-exports the small symbol surface restated in
`minpy_b200.symbol`, whose graphs are executed by hand-written sm_100a kernels
(`minpy_b200.symbol_exec`).  Nothing else of MXNet exists here; any other attribute raises.
"""
from minpy_b200 import symbol
from minpy_b200 import symbol as sym

__version__ = '0.9.2+minpy_b200_facade'


class Context(object):
    """`mx.cpu()` / `mx.gpu(i)` placeholders accepted by `Symbol.simple_bind`."""

    def __init__(self, device_type='gpu', device_id=0):
        self.device_type, self.device_id = device_type, device_id

    def __repr__(self):
        return '%s(%d)' % (self.device_type, self.device_id)


def cpu(device_id=0):
    return Context('cpu', device_id)


def gpu(device_id=0):
    return Context('gpu', device_id)


def __getattr__(name):
    raise AttributeError('mxnet façade of minpy-b200: `mxnet.%s` is not provided (only mx.sym.* / mx.symbol.*, '
                         'mx.cpu, mx.gpu); MXNet itself is not installed' % name)
