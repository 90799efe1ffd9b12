#!/usr/bin/env python
"""ORACLE tooling -- recipe that assembles `oracle/_ref/`: the UNMODIFIED reference, ready to travel.

    python oracle/build_ref.py          # build container only (needs /root/reference)

The reference is pure Python, so "building" it is a byte-for-byte copy made at build time:

    oracle/_ref/site/minpy/...            <- /root/reference/minpy            (the package, untouched)
    oracle/_ref/examples/nn/{mlp,mlp_bn_dropout,cnn,lstm}.py
                                          <- /root/reference/examples/nn/     (the four scripts of BASELINE.json)
    oracle/_ref/stubs/{mxnet,h5py}/       <- written from oracle/ref_env.py   (satisfy `import`, no arithmetic)
    oracle/_ref/MANIFEST.json             sha256 of every copied file beside its source path

`oracle/_ref/` is git-ignored (no reference source enters the history) but not gpurun-ignored, so the
GPU box gets it with the snapshot: bench.py's CPU arm then times the reference's OWN NumPy-policy path
(`cpu_baseline.kind = "reference"`) and tests/test_reference_scripts.py runs the
reference's OWN example scripts against the new `minpy`.  Nothing on the product path reads it.
"""
from __future__ import print_function

import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

SCRIPTS = ('mlp.py', 'mlp_bn_dropout.py', 'cnn.py', 'lstm.py')


def _sha(path):
    with open(path, 'rb') as f:
        return hashlib.sha256(f.read()).hexdigest()


def build(ref=None, quiet=False):
    """Assemble oracle/_ref from `ref` (default /root/reference).  Returns the directory, or None
    when the reference is not present (the GPU box: the prebuilt copy is used as it came)."""
    from oracle import ref_env
    ref = ref or os.environ.get('MINPY_REFERENCE', '/root/reference')
    dest = ref_env.REF_BUILT
    if not os.path.isdir(os.path.join(ref, 'minpy')):
        return dest if ref_env.built_prefix() else None
    if os.path.isdir(dest):
        shutil.rmtree(dest)
    manifest = {}
    site = os.path.join(dest, 'site')
    shutil.copytree(os.path.join(ref, 'minpy'), os.path.join(site, 'minpy'),
                    ignore=shutil.ignore_patterns('__pycache__', '*.pyc'))
    for base, _, files in os.walk(os.path.join(site, 'minpy')):
        for name in files:
            path = os.path.join(base, name)
            rel = os.path.relpath(path, site)
            src = os.path.join(ref, rel)
            assert _sha(src) == _sha(path), rel
            manifest[os.path.join('site', rel)] = {'source': src, 'sha256': _sha(path)}
    ex = os.path.join(dest, 'examples', 'nn')
    os.makedirs(ex)
    for name in SCRIPTS:
        src = os.path.join(ref, 'examples', 'nn', name)
        shutil.copyfile(src, os.path.join(ex, name))
        manifest[os.path.join('examples', 'nn', name)] = {'source': src, 'sha256': _sha(src)}
    ref_env.write_stubs(os.path.join(dest, 'stubs'))
    with open(os.path.join(dest, 'MANIFEST.json'), 'w') as f:
        json.dump(manifest, f, indent=1, sort_keys=True)
    if not quiet:
        print('oracle/_ref: %d files copied unmodified from %s' % (len(manifest), ref))
    return dest


if __name__ == '__main__':
    if build() is None:
        raise SystemExit('reference not found (this recipe runs in the build container only)')
