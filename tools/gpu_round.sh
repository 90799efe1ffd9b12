#!/bin/bash
# One GPU-box visit: the GPU suite, the default bench line, the reference arm, toggled variants of the
# mlp_c3 step (PDL on/off, forward fusion on/off, cluster split-K), and the ncu launch list of the default
# command.  Usage (from the repo root, through gpurun):  bash tools/gpu_round.sh <tag>
tag=${1:-run}
out=gpurun_out/$tag
mkdir -p $out
python -m pytest tests -m gpu -q > $out/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $out/pytest.log
tail -3 $out/pytest.log
quick="--no-ufunc --no-cpu-baseline --no-workloads --steps 50 --warmup 5"
for w in mlp_c3 mlp_c1 cnn_c4 lstm_c5; do
  python bench.py $quick --workload $w > $out/bench_$w.json 2> $out/bench_$w.err; echo "bench $w rc=$?"
done
for m in 0 2; do
  for w in mlp_c3 lstm_c5 cnn_c4; do
    MC_PDL=$m python bench.py $quick --no-family-replay --workload $w > $out/bench_${w}_pdl$m.json 2> $out/bench_${w}_pdl$m.err
  done
done
MINPY_FUSE_FORWARD=0 python bench.py $quick > $out/bench_mlp_c3_nofuse.json 2> $out/bench_mlp_c3_nofuse.err
for w in mlp_c3 mlp_c1 cnn_c4 lstm_c5; do
  MINPY_SIDE_STREAM=0 python bench.py $quick --no-family-replay --workload $w > $out/bench_${w}_noside.json 2> $out/bench_${w}_noside.err
done
MC_GEMM_CLUSTER=4 python bench.py $quick > $out/bench_mlp_c3_cluster4.json 2> $out/bench_mlp_c3_cluster4.err
python - <<PY
import json, glob, os
for f in sorted(glob.glob('$out/bench_*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        r = d['roofline']
        print('%-34s %8.4f ms/step  e2e %8.4f  launches/step %6.1f  top %s %.2f us x %.1f frac %.3f' % (
            os.path.basename(f), d['ms_per_step'], d['e2e']['ms_per_step'], d['launches_per_step'], r['kernel'].split()[0],
            r['avg_us_per_launch'], r['launches_per_step'], r['frac']))
    except Exception as e:
        print(os.path.basename(f), 'FAILED', e)
PY
python tools/gemm_sweep.py > $out/gemm_sweep.log 2>&1
MC_GEMM_RASTER=0 python tools/gemm_sweep.py sq4096 sq4096_nt sq8192 > $out/gemm_sweep_noraster.log 2>&1
MC_GEMM_PERSISTENT=0 python tools/gemm_sweep.py sq2048 sq4096 sq4096_nt sq4096_tn sq8192 c5_vocab > $out/gemm_sweep_nonpersistent.log 2>&1
tail -30 $out/gemm_sweep.log
python bench.py > $out/bench_default.json 2> $out/bench_default.err; echo "default bench rc=$?"
python bench.py --impl reference --steps 20 --warmup 5 > $out/bench_reference.json 2> $out/bench_reference.err; echo "reference rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file $out/launches_mlp_c3.csv \
    python bench.py $quick --steps 10 > $out/ncu_launches.log 2>&1; echo "ncu rc=$?"
