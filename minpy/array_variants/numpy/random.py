"""`minpy.numpy.random` for the NumPy variant: every routine of numpy.random, all with zero
gradient where one is ever asked for (reference: array_variants/numpy/random.py:11-18)."""
from __future__ import absolute_import

import inspect

import numpy


def register_primitives(reg, prim_wrapper):
    for name, obj in vars(numpy.random).items():
        if inspect.isroutine(obj) and not name.startswith('_'):
            reg.register(name, prim_wrapper(obj))


def def_grads(prims):
    for name in ('random', 'randn', 'rand', 'normal', 'uniform'):
        prims(name).def_grad_zero()
