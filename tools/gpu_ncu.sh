#!/bin/bash
# `ncu --set full` captures (one GPU): the GEMM family and the fused batchnorm kernels of the mlp_c3 step, the
# 4096^3 GEMM (persistent kernel), the convolution kernels of cnn_c4.  Each command has run without ncu first
# (tools/gpu_round.sh).  Raw CSV pages are written beside the reports so that reports > 64 MiB can stay behind.
tag=${1:-ncu}
out=gpurun_out/$tag
mkdir -p $out
quick="--no-ufunc --no-cpu-baseline --no-workloads --no-family-replay --steps 3 --warmup 3 --min-ms 0"
cap() { # name regex count command...
  name=$1; regex=$2; count=$3; shift 3
  ncu --set full --clock-control none --import-source on -k "regex:$regex" -c $count -f -o $out/$name "$@" > $out/$name.log 2>&1
  echo "$name rc=$?"
  ncu -i $out/$name.ncu-rep --page raw --csv > $out/${name}_raw.csv 2>/dev/null
  sz=$(stat -c %s $out/$name.ncu-rep 2>/dev/null || echo 0)
  echo "$name.ncu-rep $sz bytes"
  if [ "$sz" -gt 6000000 ]; then rm -f $out/$name.ncu-rep; fi     # gpurun_out/ merges back at most 64 MiB: keep the CSV page
}
cap prof_gemm_mlp_c3 'gemm_tf32|splitk_reduce|skinny' 30 python bench.py $quick --graph 0
cap prof_bn_mlp_c3 'bn_slab|softmax_ce_small|multi_update' 24 python bench.py $quick --graph 0
cap prof_gemm_4096 'gemm_tf32' 6 python tools/gemm_sweep.py sq4096 sq4096_nt
cap prof_gemm_8192 'gemm_tf32' 3 python tools/gemm_sweep.py sq8192
cap prof_conv_cnn_c4 'conv_igemm|conv_wgrad|maxpool' 30 python bench.py $quick --graph 0 --workload cnn_c4
cap prof_lstm_c5 'gemm_tf32|lstm_' 16 python bench.py $quick --graph 0 --workload lstm_c5
du -sh $out
