"""Two ranks on two GPUs (skipped on a one-GPU box): data-parallel step graph with the library's
NCCL all-reduce equals single-process training on the global batch (tools/dp_check.py)."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
@pytest.mark.parametrize('rule', ['sgd_momentum', 'adam'])
def test_two_rank_graph_training_equals_global_batch(rule):
    import ctypes
    from minpy_b200._lib import get_lib
    n = ctypes.c_int()
    get_lib().mc_device_count(ctypes.byref(n))
    if n.value < 2:
        pytest.skip('needs 2 GPUs')
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', '2', '--master-addr',
           '127.0.0.1', '--master-port', '29533', os.path.join(ROOT, 'tools', 'dp_check.py'), rule]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=150, cwd=ROOT)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert 'dp_check ok' in res.stdout
