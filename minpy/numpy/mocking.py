"""The mocked `numpy` namespace: a module-like object whose attributes are primitives.

Mirrors minpy/numpy/mocking.py:19-168 of the reference.  Every array variant listed in
`minpy.array_variants.variants` registers its primitives and VJPs into one `Registry`;
`generate_attrs(policy)` then binds each name to the primitive the policy picks (static
dispatch), or to a `PrimitiveSelector` when op statistics are being recorded.  Unlike the
reference, names the policy rejects are removed again when the policy changes, so that
`only_cuda` really exposes nothing but CUDA primitives.
"""
from __future__ import absolute_import

import importlib

import minpy
from minpy.array_variants import variants
from minpy.dispatch.primitive_selector import PrimitiveSelector
from minpy.dispatch.registry import Registry
from minpy.primitive import Primitive
from minpy.utils import log


class Module(object):
    """Stands in for a numpy (sub)module; see the module docstring."""

    def __init__(self, old, name=None, name_injector=None):
        minpy.Config['modules'].append(self)
        self._registry = Registry(old['__name__'])
        self._logger = log.get_logger(old['__name__'])
        self._old = old
        self._name_injector = name_injector if name_injector else {}
        self._bound = set()
        for vname, vtype in variants.items():
            modname = 'minpy.array_variants.' + vname + ('' if name is None else '.' + name)
            mod = importlib.import_module(modname)
            mod.register_primitives(
                self._registry,
                (lambda ty: lambda func, *a, **kw: Primitive(func, ty, *a, **kw))(vtype))
            mod.def_grads((lambda ty: lambda pname: self._registry.get(pname, ty))(vtype))
        self.generate_attrs(minpy.Config['default_policy'])

    def set_policy(self, policy):
        """Deprecated spelling of `generate_attrs`."""
        self.generate_attrs(policy)

    def record_op_stat(self):
        """Count every dispatched call from now on (per-call dispatch through the policy)."""
        self.generate_attrs(self.policy, True)

    def show_op_stat(self):
        self.policy.show_op_stat()

    def generate_attrs(self, policy, use_selector=False):
        """(Re)bind the namespace for `policy`; later sources override earlier ones:
        original module attributes < injected NumPy constants/types < primitives."""
        for stale in self._bound:
            if stale in self.__dict__:
                delattr(self, stale)
        self._bound = set()
        self.policy = policy
        for key, val in self._old.items():
            setattr(self, key, val)
        for key in self._name_injector.keys():
            setattr(self, key, self._name_injector[key])
        for key, prims in self._registry._reg.items():  # pylint: disable=protected-access
            if use_selector:
                setattr(self, key, PrimitiveSelector(key, self))
                self._bound.add(key)
            else:
                choice = policy.decide(prims.values())
                if choice is not None:
                    setattr(self, key, prims[choice])
                    self._bound.add(key)
        self.__registry__ = self._registry


class NameInjector(object):
    """Read-only name -> object table copied from a real module: objects whose type is in
    `injected_type` (constants, dtypes), names in `name_set`, and `exception` renames."""

    def __init__(self, dest_mod, name='', injected_type=None, name_set=None, exception=None):
        self.__name__ = '<%s>' % name if name else name
        source = vars(dest_mod)
        table = {}
        if injected_type is not None:
            kinds = tuple(injected_type)
            for key, obj in source.items():
                if type(obj) in kinds:  # pylint: disable=unidiomatic-typecheck
                    table[key] = obj
        for key in (name_set or ()):
            table[key] = source[key]
        for key, target in (exception or {}).items():
            table[key] = source[target]
        self._name_dict = table

    def keys(self):
        return self._name_dict.keys()

    def __len__(self):
        return len(self._name_dict)

    def __contains__(self, key):
        return key in self._name_dict

    def __getitem__(self, name):
        return self._name_dict[name]
