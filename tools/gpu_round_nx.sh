#!/bin/bash
# N-GPU visit (N = 4 or 8): peer-memory all-reduce check + timings against NCCL, and the bench line with the
# window forced on / off.  Usage: gpurun --gpus N -- bash tools/gpu_round_nx.sh <tag> <N>
tag=${1:-nx}; NP=${2:-4}
out=gpurun_out/$tag
mkdir -p $out
tr() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $NP --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
MINPY_P2P_ALLREDUCE=1 tr 29611 tools/p2p_check.py > $out/p2p_check.log 2>&1; echo "p2p_check rc=$?"; tail -2 $out/p2p_check.log
MINPY_P2P_ALLREDUCE=1 tr 29612 tools/comm_bench.py > $out/comm_bench.log 2>&1; grep "^world" $out/comm_bench.log
bench() { tr $1 bench.py --gpus $NP --steps 50 --warmup 5 "${@:2}"; }
MINPY_P2P_ALLREDUCE=1 bench 29613 > $out/bench_p2p.json 2> $out/bench_p2p.err; echo "bench p2p rc=$?"
MINPY_P2P_ALLREDUCE=0 bench 29614 --no-workloads > $out/bench_nccl.json 2> $out/bench_nccl.err; echo "bench nccl rc=$?"
python - <<PY
import json, glob, os
for f in sorted(glob.glob('$out/bench_*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print('%-22s n=%d %10.0f samples/s %8.4f ms/step e2e %8.4f  dp_parity %s' % (os.path.basename(f), d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['ms_per_step'], (d.get('dp_parity') or {}).get('ok')))
        for w, v in d.get('workloads', {}).items():
            print('    %-10s %10.0f samples/s %8.4f ms/step' % (w, v['value'], v['ms_per_step']))
    except Exception as e:
        print(os.path.basename(f), 'FAILED', e)
PY
