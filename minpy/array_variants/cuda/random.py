"""`minpy.numpy.random` for the CUDA variant: Philox4x32-10 streams generated on the device.

Replaces the host RNG + host->device copy of the reference (`random.rand` was NumPy-only,
docs/feature/op_stat.rst:114; mx.random.uniform/normal in array_variants/mxnet/random.py:9-16).
The stream is keyed by (seed, offset); every draw advances the offset, `seed()` resets it.
Ranks of a data-parallel job fold their rank into the seed so that dropout masks differ.
"""
from __future__ import absolute_import

import os

import numpy as np

from minpy_b200 import ops
from minpy_b200.cuda_array import CudaArray

_state = {'seed': 0x5EED0000 + int(os.environ.get('RANK', '0')), 'offset': 0}


def next_stream(n):
    """Reserve `n` draws; returns the (seed, offset) that generates them."""
    seed, offset = _state['seed'], _state['offset']
    _state['offset'] += (int(n) + 3) // 4 + 1
    return seed, offset


def seed(value=None):
    _state['seed'] = (int(value) if value is not None else int.from_bytes(os.urandom(4), 'little')) & 0xFFFFFFFFFFFFFFFF
    _state['offset'] = 0


def _shape_of(size):
    if size is None:
        return ()
    if isinstance(size, (int, np.integer)):
        return (int(size),)
    return tuple(int(s) for s in size)


def _draw(kind, shape):
    out = CudaArray.empty(shape)
    if out.size:
        sd, off = next_stream(out.size)
        getattr(ops.L(), kind)(out.ptr, out.size, sd, off)
    return out


def rand(*shape):
    return _draw('mc_rand_uniform', tuple(int(s) for s in shape))


def randn(*shape):
    return _draw('mc_rand_normal', tuple(int(s) for s in shape))


def random(size=None):
    return _draw('mc_rand_uniform', _shape_of(size))


random_sample = random


def uniform(low=0.0, high=1.0, size=None):
    u = _draw('mc_rand_uniform', _shape_of(size))
    lib = ops.L()
    return ops.binary(lib.MC_ADD, ops.binary(lib.MC_MUL, u, float(high) - float(low)), float(low))


def normal(loc=0.0, scale=1.0, size=None):
    z = _draw('mc_rand_normal', _shape_of(size))
    lib = ops.L()
    return ops.binary(lib.MC_ADD, ops.binary(lib.MC_MUL, z, float(scale)), float(loc))


def randint(low, high=None, size=None):
    if high is None:
        low, high = 0, low
    u = _draw('mc_rand_uniform', _shape_of(size))
    lib = ops.L()
    scaled = ops.unary(lib.MC_FLOOR, ops.binary(lib.MC_ADD, ops.binary(lib.MC_MUL, u, float(high - low)), float(low)))
    return scaled.astype(np.int64)


def register_primitives(reg, prim_wrapper):
    for name, fn in (('rand', rand), ('randn', randn), ('random', random), ('random_sample', random),
                     ('uniform', uniform), ('normal', normal), ('randint', randint), ('seed', seed)):
        reg.register(name, prim_wrapper(fn))


def def_grads(prims):
    for name in ('rand', 'randn', 'random', 'uniform', 'normal'):
        prims(name).def_grad_zero()
