"""`import minpy.numpy as np`: replaced in sys.modules by a mocked namespace whose functions
are MinPy primitives and whose constants / dtypes are NumPy's own."""
import sys

import numpy

from minpy.array import Value
from minpy.numpy.mocking import Module, NameInjector
from . import random

_mod = Module(
    {'__path__': __path__, '__name__': __name__, 'random': random},
    name_injector=NameInjector(numpy, name='numpy injector',
                               injected_type={float, int, type(None), type}))
Value._ns = _mod  # pylint: disable=protected-access
sys.modules[__name__] = _mod
