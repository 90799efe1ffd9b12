#!/bin/bash
# Two-GPU visit: the 2-GPU tests (data-parallel parity, peer-memory all-reduce, `with gpu(1):` placement) and
# the N = 2 bench line with dp_parity.  Usage: gpurun --gpus 2 -- bash tools/gpu_round_n2.sh <tag>
tag=${1:-n2}
out=gpurun_out/$tag
mkdir -p $out
python -m pytest tests/test_parallel_gpu.py tests/test_rows.py tests/test_context_logic.py -m gpu -q > $out/pytest_n2.log 2>&1; echo "pytest rc=$?" | tee -a $out/pytest_n2.log
tail -3 $out/pytest_n2.log
tr() { python -m torch.distributed.run --nnodes=1 --nproc-per-node ${NP:-2} --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
for push in 1 0; do
  MINPY_P2P_PUSH=$push tr 2952$push tools/p2p_check.py > $out/p2p_check_push$push.log 2>&1; echo "p2p_check push=$push rc=$?"; tail -2 $out/p2p_check_push$push.log
  MINPY_P2P_PUSH=$push tr 2953$push tools/comm_bench.py > $out/comm_bench_push$push.log 2>&1; grep -v "^\[" $out/comm_bench_push$push.log | tail -10
done
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 bench.py --gpus 2 --steps 50 --warmup 5 "${@:2}"; }
run 29511 > $out/bench_n2.json 2> $out/bench_n2.err; echo "bench n2 rc=$?"
MINPY_SIDE_STREAM=0 run 29512 --no-workloads > $out/bench_n2_noside.json 2> $out/bench_n2_noside.err; echo "bench n2 noside rc=$?"
MINPY_P2P_PUSH=0 run 29513 --no-workloads > $out/bench_n2_pull.json 2> $out/bench_n2_pull.err; echo "bench n2 pull rc=$?"
MINPY_P2P_ALLREDUCE=0 run 29514 --no-workloads > $out/bench_n2_nccl.json 2> $out/bench_n2_nccl.err; echo "bench n2 nccl rc=$?"
python bench.py --no-ufunc --no-cpu-baseline --no-workloads --steps 50 --warmup 5 > $out/bench_n1.json 2> $out/bench_n1.err
python - <<PY
import json, glob, os
for f in sorted(glob.glob('$out/bench_*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print('%-28s n=%d %10.0f samples/s %8.4f ms/step e2e %8.4f  dp_parity %s' % (os.path.basename(f), d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['ms_per_step'], d.get('dp_parity')))
        for w, v in d.get('workloads', {}).items():
            print('    %-10s %10.0f samples/s %8.4f ms/step' % (w, v['value'], v['ms_per_step']))
    except Exception as e:
        print(os.path.basename(f), 'FAILED', e)
PY
