"""Registry tables of the array variants (the plug-in point of the reference,
minpy/array_variants/__init__.py:12-21).

Two variants ship here: `numpy` (host; the reference's own oracle variant) and `cuda`
(libminpy_cuda.so, sm_100a).  `ArrayType.MXNET` keeps its number so that code comparing
against it still imports, but no MXNet variant is registered.
"""
from __future__ import absolute_import

from minpy.array_variants import numpy
from minpy.array_variants import cuda


class ArrayType(object):
    """Enumeration of the array back ends."""
    NUMPY = 0
    MXNET = 1   # reserved: the reference's second variant, replaced by CUDA
    CUDA = 2


variants = {'numpy': ArrayType.NUMPY, 'cuda': ArrayType.CUDA}
array_types = {'numpy': numpy.array_type, 'cuda': cuda.array_type}
variants_repr = {ArrayType.NUMPY: 'NumPy', ArrayType.CUDA: 'CUDA'}
number_types = {'native': [int, float], 'numpy': numpy.number_type, 'cuda': cuda.number_type}
