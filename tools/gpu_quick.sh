#!/bin/bash
# Short GPU-box visit: the GPU suite and one bench line per workload (no variants, no ncu).
# Usage (from the repo root, through gpurun):  bash tools/gpu_quick.sh <tag>
tag=${1:-quick}
out=gpurun_out/$tag
mkdir -p $out
python -m pytest tests -m gpu -q -x > $out/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $out/pytest.log
tail -3 $out/pytest.log
quick="--no-ufunc --no-cpu-baseline --no-workloads --steps 50 --warmup 5"
for w in mlp_c3 mlp_c1 cnn_c4 lstm_c5; do
  python bench.py $quick --workload $w > $out/bench_$w.json 2> $out/bench_$w.err; echo "bench $w rc=$?"
  grep -h "replay" $out/bench_$w.err | head -8
done
python - <<PY
import json, glob, os
for f in sorted(glob.glob('$out/bench_*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print('%-28s %8.4f ms/step  e2e %8.4f  launches/step %6.1f' % (os.path.basename(f), d['ms_per_step'], d['e2e']['ms_per_step'], d['launches_per_step']))
    except Exception as e:
        print(os.path.basename(f), 'FAILED', e)
PY
