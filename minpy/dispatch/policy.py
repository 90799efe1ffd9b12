"""Policies choosing which array variant serves a name (reference:
minpy/dispatch/policy.py:36-131,220-301), generalised from the hard-wired NumPy/MXNet pair
to any set of variants.  `prefer_cuda` / `only_cuda` take the place of `prefer_mxnet` /
`only_mxnet`; the old names are accepted as aliases so that existing scripts keep working.
"""
from __future__ import absolute_import
from __future__ import print_function

import functools
from collections import defaultdict

import minpy
from minpy import tape
from minpy.array_variants import ArrayType, variants_repr
from minpy.utils import log

_logger = log.get_logger(__name__)


class PrimitivePolicyError(ValueError):
    """No implementation of a name satisfies the policy."""

    def __init__(self, name, policy_name):
        super(PrimitivePolicyError, self).__init__(
            'Cannot find implementation for function: {}() under policy: {}. '
            'Maybe lack of gradient implementation?'.format(name, policy_name))


class Policy(object):
    """Base class: `decide` ranks variants, `resolve_call` performs a counted call."""
    order = ()   # variants in decreasing preference

    def __init__(self):
        self._op_stat = defaultdict(lambda: defaultdict(int))

    # the reference exposed these two counters by name
    @property
    def _numpy_op_stat(self):
        return self._op_stat[ArrayType.NUMPY]

    @property
    def _cuda_op_stat(self):
        return self._op_stat[ArrayType.CUDA]

    @property
    def name(self):
        return type(self).__name__

    def decide(self, candidates):
        """Variant to use among `candidates` (primitives of one name), or None."""
        present = set(p.type for p in candidates)
        for vtype in self.order:
            if vtype in present:
                return vtype
        return None

    def resolve_call(self, name, reg, args, kwargs):
        current = tape.global_tape()
        bp_args = tuple(i for i, a in enumerate(args)
                        if hasattr(a, 'is_marked_for_bp') and a.is_marked_for_bp(current))
        bp_kwargs = tuple(k for k, a in kwargs.items()
                          if hasattr(a, 'is_marked_for_bp') and a.is_marked_for_bp(current))
        choice = self.decide(reg.iter_available_types(name, bp_args, bp_kwargs))
        if choice is None:
            raise PrimitivePolicyError(name, self.name)
        self._op_stat[choice][name] += 1
        return reg.get(name, choice).call(args, kwargs)

    def show_op_stat(self):
        """Print how many calls each variant served."""
        print('--------Op Dispatch Statistics Start--------')
        totals = {}
        for vtype in sorted(variants_repr):
            print('%s op called times:' % variants_repr[vtype].upper())
            totals[vtype] = 0
            for op, count in sorted(self._op_stat[vtype].items()):
                print(' {} : {}'.format(op, count))
                totals[vtype] += count
        grand = sum(totals.values())
        if grand:
            print('Total Dispatch Proportion: ' + ', '.join(
                '{:.1%} in {}'.format(totals[v] / float(grand), variants_repr[v]) for v in sorted(totals)))
        print('--------Op Dispatch Statistics End--------')


class PreferCudaPolicy(Policy):
    """CUDA where a kernel exists, NumPy (with a device->host copy) for the long tail."""
    order = (ArrayType.CUDA, ArrayType.NUMPY)


class OnlyNumPyPolicy(Policy):
    order = (ArrayType.NUMPY,)


class OnlyCudaPolicy(Policy):
    """Never leaves the device: names without a CUDA primitive are simply absent."""
    order = (ArrayType.CUDA,)


# the reference's class names, for scripts that import them
PreferMXNetPolicy = PreferCudaPolicy
OnlyMXNetPolicy = OnlyCudaPolicy

_BY_NAME = {
    'prefer_cuda': PreferCudaPolicy, 'only_cuda': OnlyCudaPolicy, 'only_numpy': OnlyNumPyPolicy,
    'prefer_mxnet': PreferCudaPolicy, 'only_mxnet': OnlyCudaPolicy,
}


def create(plc):
    """Policy object from a name or an existing policy."""
    if isinstance(plc, Policy):
        return plc
    try:
        return _BY_NAME[plc]()
    except (KeyError, TypeError):
        raise TypeError('Unknown policy name %s' % (plc,))


def wrap_policy(plc):
    """Decorator running a function under `plc` and restoring the previous policy after."""
    def decorator(func):
        @functools.wraps(func)
        def wrapper(*args, **kwargs):
            previous = minpy.get_global_policy()
            minpy.set_global_policy(plc)
            try:
                return func(*args, **kwargs)
            finally:
                minpy.set_global_policy(previous)
        return wrapper
    return decorator
