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


def _two_gpus():
    import ctypes
    from minpy_b200._lib import get_lib
    n = ctypes.c_int()
    get_lib().mc_device_count(ctypes.byref(n))
    return n.value >= 2


def _torchrun(script, args, port, env=None):
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', '2', '--master-addr',
           '127.0.0.1', '--master-port', str(port), os.path.join(ROOT, 'tools', script)] + list(args)
    full_env = dict(os.environ)
    full_env.update(env or {})
    return subprocess.run(cmd, capture_output=True, text=True, timeout=150, cwd=ROOT, env=full_env)


@pytest.mark.gpu
def test_peer_memory_allreduce_is_exact_and_identical_on_all_ranks():
    """mc_allreduce_avg inside the IPC window (two-shot all-reduce over NVLink peer loads): equal to the
    float32 mean summed in rank order, bitwise the same on both ranks, eager and in graph replays."""
    if not _two_gpus():
        pytest.skip('needs 2 GPUs')
    res = _torchrun('p2p_check.py', [], 29534)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert 'p2p_check ok' in res.stdout


@pytest.mark.gpu
def test_two_rank_training_through_nccl_when_the_window_is_off():
    """MINPY_P2P_ALLREDUCE=0 keeps the NCCL all-reduce: the same parity as above holds on that path."""
    if not _two_gpus():
        pytest.skip('needs 2 GPUs')
    res = _torchrun('dp_check.py', ['sgd_momentum'], 29535, env={'MINPY_P2P_ALLREDUCE': '0'})
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert 'dp_check ok' in res.stdout
