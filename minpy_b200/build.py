"""In-tree build of libminpy_cuda.so for sm_100a (nvcc cross-compiles without a GPU).

Usage: ``python -m minpy_b200.build [--force] [--verbose]``.  The shared library is
written next to this file so that it travels with the repository snapshot to the GPU box;
it is git-ignored (built artefact).
"""
from __future__ import annotations

import concurrent.futures
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
BUILD_DIR = os.path.join(HERE, '_build')
LIB_PATH = os.path.join(HERE, 'libminpy_cuda.so')
HEADER = os.path.join(os.path.dirname(HERE), 'include', 'minpy_cuda.h')

SOURCES = ['runtime.cu', 'elementwise.cu', 'reduce.cu', 'softmax.cu', 'index.cu', 'gemm.cu', 'layers.cu', 'lstm.cu', 'lstm_persist.cu', 'conv.cu', 'conv_tc.cu', 'comm.cu']

NVCC_FLAGS = [
    '-gencode', 'arch=compute_100a,code=sm_100a',
    '-O3', '-std=c++17', '-lineinfo',
    '-Xcompiler', '-fPIC',
    '--expt-relaxed-constexpr',
    '-Xptxas', '-v',
]


def _nvcc() -> str:
    exe = shutil.which('nvcc') or '/usr/local/cuda/bin/nvcc'
    if not os.path.exists(exe):
        raise RuntimeError('nvcc not found; libminpy_cuda.so cannot be built')
    return exe


def _stale(target: str, deps: list[str]) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _compile(src: str, force: bool, verbose: bool) -> str:
    obj = os.path.join(BUILD_DIR, os.path.splitext(src)[0] + '.o')
    deps = [os.path.join(CSRC, src), os.path.join(CSRC, 'common.cuh'), os.path.join(CSRC, 'tc_common.cuh'), HEADER, os.path.abspath(__file__)]
    if force or _stale(obj, deps):
        cmd = [_nvcc()] + NVCC_FLAGS + ['-c', os.path.join(CSRC, src), '-o', obj]
        res = subprocess.run(cmd, capture_output=True, text=True)
        log = os.path.join(BUILD_DIR, os.path.splitext(src)[0] + '.log')
        with open(log, 'w') as f:
            f.write(' '.join(cmd) + '\n' + res.stdout + res.stderr)
        if res.returncode != 0:
            raise RuntimeError('nvcc failed for %s:\n%s' % (src, res.stderr[-6000:]))
        if verbose:
            print(res.stderr, file=sys.stderr)
    return obj


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile every CUDA source and link the shared library; returns its path."""
    os.makedirs(BUILD_DIR, exist_ok=True)
    with concurrent.futures.ThreadPoolExecutor(max_workers=min(8, len(SOURCES))) as pool:
        objs = list(pool.map(lambda s: _compile(s, force, verbose), SOURCES))
    if force or _stale(LIB_PATH, objs):
        # export only the C ABI (mc_*)
        vers = os.path.join(BUILD_DIR, 'exports.map')
        with open(vers, 'w') as f:
            f.write('{ global: mc_*; local: *; };\n')
        cmd = [_nvcc(), '-shared', '-o', LIB_PATH] + objs + [
            '-gencode', 'arch=compute_100a,code=sm_100a', '-Xlinker', '--version-script=' + vers, '-ldl']
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError('link failed:\n' + res.stderr[-4000:])
    return LIB_PATH


if __name__ == '__main__':
    path = build(force='--force' in sys.argv, verbose='--verbose' in sys.argv)
    print(path)
