"""`minpy.numpy.random`: replaced in sys.modules by a mocked namespace."""
from __future__ import absolute_import

import sys

from minpy.numpy.mocking import Module

sys.modules[__name__] = Module({'__name__': __name__}, 'random')
