"""pytest configuration: registers the `gpu` marker and puts the repo root on sys.path."""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box)')


def has_gpu():
    try:
        import ctypes
        from minpy_b200._lib import get_lib
        n = ctypes.c_int(0)
        get_lib().mc_device_count(ctypes.byref(n))
        return n.value > 0
    except Exception:
        return False


@pytest.fixture(scope='session')
def cuda():
    """The functional CUDA ops module, initialised on device LOCAL_RANK / 0."""
    from minpy_b200 import ops
    ops.L()
    return ops


# ---- measured parity errors: every tolerance in the suite should sit ~3x above what is measured ----
PARITY_ERRORS = {}


def record_err(key, value):
    """Keep the worst measured error per key; written to gpurun_out/parity_errors.json at session end."""
    value = float(value)
    if value > PARITY_ERRORS.get(key, -1.0):
        PARITY_ERRORS[key] = value
    return value


def pytest_sessionfinish(session, exitstatus):  # pylint: disable=unused-argument
    if not PARITY_ERRORS:
        return
    import json
    out = os.path.join(ROOT, 'gpurun_out')
    try:
        os.makedirs(out, exist_ok=True)
        with open(os.path.join(out, 'parity_errors.json'), 'w') as f:
            json.dump(PARITY_ERRORS, f, indent=1, sort_keys=True)
    except OSError:
        pass
