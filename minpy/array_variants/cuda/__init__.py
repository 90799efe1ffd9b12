"""CUDA array variant: device arrays (float32 / int32 / int64) served by libminpy_cuda.so.

Plays the role the MXNet variant played in the reference
(minpy/array_variants/mxnet/__init__.py:10-14) behind the same four attributes.
"""
from __future__ import absolute_import

from minpy_b200.cuda_array import CudaArray
from minpy.array_variants.cuda import cuda_core

array_type = CudaArray
number_type = []

register_primitives = cuda_core.register_primitives
def_grads = cuda_core.def_grads
