#!/bin/bash
# Build tools/_probe/libminpy_cuda_probe.so: the product objects, with the sources named on the command line
# (default: softmax.cu layers.cu) recompiled under -DMC_PROBE so that their MC_STAMP points store %globaltimer.
# Run here, before the gpurun call; tools/probe.py reads the stamps on the box:
#   MINPY_CUDA_LIB=$PWD/tools/_probe/libminpy_cuda_probe.so python tools/probe.py head 1024 512 10
set -e
cd "$(dirname "$0")/.."
python -m minpy_b200.build > /dev/null
mkdir -p tools/_probe
srcs=${@:-softmax.cu layers.cu}
objs=$(ls minpy_b200/_build/*.o)
for s in $srcs; do
  b=${s%.cu}
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC --expt-relaxed-constexpr \
    -DMC_PROBE -c minpy_b200/csrc/$s -o tools/_probe/${b}_probe.o
  objs=$(echo "$objs" | grep -v "/${b}.o") 
  objs="$objs tools/_probe/${b}_probe.o"
done
nvcc -shared -o tools/_probe/libminpy_cuda_probe.so $objs -gencode arch=compute_100a,code=sm_100a -ldl
ls -la tools/_probe/libminpy_cuda_probe.so
