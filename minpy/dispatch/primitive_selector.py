"""Per-call dispatch thunk installed by `record_op_stat()` (reference:
minpy/dispatch/primitive_selector.py:8-31): asks the module's policy on every call, so that
gradient availability is considered and the call is counted."""
from __future__ import absolute_import


class PrimitiveSelector(object):
    __slots__ = ['_name', '_mod']

    def __init__(self, name, mod):
        self._name = name
        self._mod = mod

    @property
    def name(self):
        return self._name

    def __call__(self, *args, **kwargs):
        return self._mod.policy.resolve_call(self._name, self._mod.__registry__, args, kwargs)
