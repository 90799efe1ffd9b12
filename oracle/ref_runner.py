#!/usr/bin/env python
"""ORACLE tooling -- times the UNMODIFIED reference's NumPy-policy training step (the CPU baseline).

    python oracle/ref_runner.py --workload mlp_c3 --steps 20 --warmup 3 [--budget 25] [--threads N]

Runs in its OWN process (the reference's package is also called `minpy`): oracle/ref_env.activate puts
oracle/_ref/site (the copy oracle/build_ref.py made) first on sys.path, sets `only_numpy`, then the
workload's model -- the same `ModelBase` subclass bench.py trains on the GPU (minpy_b200/workloads.py,
written against the reference's `nn.layers` / `nn.model` API) -- is trained by the reference's own
`Solver._step` (nn/solver.py:157-195: `core.grad_and_loss`, the reference tape, its `optim` rule), in
float64 as its NumPy variant computes, on all host cores BLAS will use.  Prints ONE JSON line.

Only bench.py's CPU legs and tests/ execute this file; it is never on the product path.
"""
from __future__ import print_function

import argparse
import json
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--workload', default='mlp_c3')
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--budget', type=float, default=25.0, help='stop after this many seconds of timed steps')
    ap.add_argument('--threads', type=int, default=0, help='BLAS threads (0 = every host core)')
    ap.add_argument('--prefix', default=None, help='directory holding the reference `minpy` (default oracle/_ref/site)')
    args = ap.parse_args()

    threads = args.threads or os.cpu_count() or 1
    # torchrun exports OMP_NUM_THREADS=1 to its workers: the baseline must not inherit that
    for var in ('OMP_NUM_THREADS', 'OPENBLAS_NUM_THREADS', 'MKL_NUM_THREADS'):
        os.environ[var] = str(threads)
    if ROOT not in sys.path:
        sys.path.insert(0, ROOT)
    from oracle import ref_env
    minpy = ref_env.activate(args.prefix, keep_root=True)
    import numpy as onp
    from threadpoolctl import threadpool_limits, threadpool_info
    from minpy.nn.solver import Solver
    from minpy.nn.io import DataBatch
    import minpy.numpy as mnp
    from minpy_b200 import workloads            # models written against the reference's nn API

    factory, batch, rule, optim_config, make_data, desc = workloads.WORKLOADS[args.workload]
    model = factory()
    solver = Solver(model, None, None, update_rule=rule, optim_config=dict(optim_config), init_rule='gaussian',
                    init_config={'stdvar': 0.01}, verbose=False)
    solver.init()
    for k, v in workloads.init_params(model, 3).items():       # the same initial weights as every other arm
        model.params[k] = mnp.array(v.astype(onp.float64))
    X, Y = make_data(batch * 4, seed=100)
    batches = [DataBatch([mnp.array(X[j * batch:(j + 1) * batch].astype(onp.float64))],
                         [mnp.array(Y[j * batch:(j + 1) * batch])], 0, j) for j in range(4)]
    with threadpool_limits(limits=threads):
        for i in range(args.warmup):
            solver._step(batches[i % 4])  # pylint: disable=protected-access
        t0 = time.perf_counter()
        n = 0
        while n < args.steps:
            solver._step(batches[n % 4])  # pylint: disable=protected-access
            n += 1
            if time.perf_counter() - t0 > args.budget:
                break
        dt = time.perf_counter() - t0
        info = threadpool_info()
    loss = float(onp.asarray(solver.loss_history[-1]))
    print(json.dumps({
        'workload': args.workload, 'description': desc, 'batch': batch, 'steps': n, 'warmup': args.warmup,
        'seconds': dt, 'ms_per_step': 1e3 * dt / n, 'value': batch * n / dt, 'unit': 'samples/s',
        'cores': int(max([p.get('num_threads', 1) for p in info] or [threads])), 'host_cpus': os.cpu_count(),
        'blas': sorted(set('%s %s' % (p.get('internal_api'), p.get('version')) for p in info)),
        'numpy': onp.__version__, 'dtype': 'f64', 'final_loss': loss, 'policy': minpy.get_global_policy().name
        if hasattr(minpy, 'get_global_policy') else 'only_numpy',
        'minpy_file': os.path.relpath(minpy.__file__, ROOT)}), flush=True)


if __name__ == '__main__':
    main()
