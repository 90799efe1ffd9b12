#!/bin/bash
# Build tools/_probe/libminpy_cuda_probe.so: the product objects with softmax.cu recompiled under -DMC_HEAD_PROBE
# (thread 0 of every block of head_softmax_kernel stamps %globaltimer at its phase boundaries).  Run here, before
# the gpurun call; tools/head_probe.py reads the stamps on the box.
set -e
cd "$(dirname "$0")/.."
python -m minpy_b200.build > /dev/null
mkdir -p tools/_probe
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC --expt-relaxed-constexpr \
  -DMC_HEAD_PROBE -c minpy_b200/csrc/softmax.cu -o tools/_probe/softmax_probe.o
objs=$(ls minpy_b200/_build/*.o | grep -v '/softmax.o')
nvcc -shared -o tools/_probe/libminpy_cuda_probe.so $objs tools/_probe/softmax_probe.o \
  -gencode arch=compute_100a,code=sm_100a -ldl
ls -la tools/_probe/libminpy_cuda_probe.so
