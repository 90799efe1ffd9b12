"""name -> {ArrayType -> Primitive} table of one mocked namespace
(reference: minpy/dispatch/registry.py:14-94)."""
from __future__ import absolute_import

from minpy.utils import log

_logger = log.get_logger(__name__)


class PrimitiveRegistryError(ValueError):
    """Raised for malformed registrations."""


class Registry(object):
    """Primitives of one namespace, grouped by name, one entry per array variant."""

    def __init__(self, namespace):
        self._reg = {}
        self._ns = namespace

    @property
    def nspace(self):
        return self._ns

    def register(self, name, prim):
        """Add `prim` under `name`; a second primitive of the same variant is ignored with a
        warning (first registration wins, as in the reference :45-47)."""
        slot = self._reg.setdefault(name, {})
        if prim.type in slot:
            _logger.warning('%s already has a %s implementation; keeping the first', name, prim.typestr)
            return
        slot[prim.type] = prim

    def has_name(self, name):
        return name in self._reg

    def exists(self, name, ptype):
        return ptype in self._reg.get(name, ())

    def get(self, name, ptype):
        return self._reg[name][ptype]

    def iter_available_types(self, name, bp_args, bp_kwargs):
        """Primitives of `name` able to differentiate w.r.t. the given argument positions /
        keywords."""
        return [p for p in self._reg.get(name, {}).values() if p.gradable(bp_args, bp_kwargs)]
