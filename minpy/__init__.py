"""MinPy front end on the B200-native `cuda` array variant.

`import minpy.numpy as np`, `minpy.core.grad / grad_and_loss`, `minpy.nn.*` keep the
reference's API (dmlc/minpy @2e44927, minpy/__init__.py:1-53); the MXNet dependency is gone:
the second array variant is `cuda` (libminpy_cuda.so, sm_100a) and the default policy
prefers it.  There is no silent CPU execution of a CUDA primitive: if the library or a
device is missing the first CUDA call raises.
"""
from __future__ import absolute_import

Config = {'modules': []}

from .dispatch import policy  # noqa: E402  pylint: disable=wrong-import-position

Config['default_policy'] = policy.PreferCudaPolicy()

from . import numpy  # noqa: E402,F401  pylint: disable=wrong-import-position


def set_global_policy(plc):
    """Switch every mocked namespace (and future ones) to policy `plc` (name or object):
    'prefer_cuda' (default), 'only_cuda', 'only_numpy'."""
    Config['default_policy'] = policy.create(plc)
    for mod in Config['modules']:
        mod.generate_attrs(Config['default_policy'])


def get_global_policy():
    return Config['default_policy']


wrap_policy = policy.wrap_policy

__version__ = '0.3.4+b200'
