"""ctypes binding of libminpy_cuda.so (the C ABI declared in include/minpy_cuda.h).

The prototypes are read from the header itself, so the binding cannot drift from the
declared ABI.  There is NO fallback: if the shared library is missing or fails to load,
importing the CUDA path raises -- the product never computes on the CPU.
"""
from __future__ import annotations

import ctypes
import os
import re

HERE = os.path.dirname(os.path.abspath(__file__))
HEADER = os.path.join(os.path.dirname(HERE), 'include', 'minpy_cuda.h')
LIB_PATH = os.environ.get('MINPY_CUDA_LIB', os.path.join(HERE, 'libminpy_cuda.so'))

_SCALAR_TYPES = {
    'int': ctypes.c_int,
    'int32_t': ctypes.c_int32,
    'int64_t': ctypes.c_int64,
    'uint64_t': ctypes.c_uint64,
    'float': ctypes.c_float,
    'size_t': ctypes.c_size_t,
}


class CudaLibError(RuntimeError):
    """A call into libminpy_cuda.so returned a non-zero status."""


def parse_header(path: str = HEADER):
    """Return {name: (restype, [argtypes], [argnames])} for every ``mc_*`` declaration."""
    text = open(path).read()
    text = re.sub(r'/\*.*?\*/', ' ', text, flags=re.S)
    text = re.sub(r'//[^\n]*', ' ', text)
    text = re.sub(r'^\s*#.*$', ' ', text, flags=re.M)
    decls = {}
    for m in re.finditer(r'\b(int|const char\s*\*)\s+(mc_\w+)\s*\(([^)]*)\)\s*;', text):
        ret, name, args = m.group(1), m.group(2), m.group(3).strip()
        argtypes, argnames = [], []
        if args and args != 'void':
            for a in args.split(','):
                a = ' '.join(a.split())
                mm = re.match(r'^(.*?)(\w+)$', a)
                ty, nm = mm.group(1).strip(), mm.group(2)
                if '*' in ty:
                    argtypes.append(ctypes.c_void_p)
                else:
                    argtypes.append(_SCALAR_TYPES[ty.replace('const ', '').strip()])
                argnames.append(nm)
        restype = ctypes.c_int if ret == 'int' else ctypes.c_char_p
        decls[name] = (restype, argtypes, argnames)
    return decls


def parse_enums(path: str = HEADER):
    """Return {enumerator: value} for every enum in the header."""
    text = open(path).read()
    text = re.sub(r'/\*.*?\*/', ' ', text, flags=re.S)
    out = {}
    for m in re.finditer(r'enum\s+\w+\s*\{([^}]*)\}', text):
        val = -1
        for item in m.group(1).split(','):
            item = item.strip()
            if not item:
                continue
            if '=' in item:
                k, v = item.split('=')
                val = int(v.strip(), 0)
                out[k.strip()] = val
            else:
                val += 1
                out[item] = val
    return out


class VjpStep(ctypes.Structure):
    """Mirror of ``mc_vjp_step``."""
    _fields_ = [('op', ctypes.c_int32), ('a_mode', ctypes.c_int32), ('b_mode', ctypes.c_int32),
                ('scalar', ctypes.c_float), ('a', ctypes.c_void_p), ('b', ctypes.c_void_p),
                ('inner', ctypes.c_int64)]


class _Lib(object):
    """Loaded library with checked entry points: ``lib.mc_foo(...)`` raises on failure."""

    def __init__(self, path: str = LIB_PATH):
        if not os.path.exists(path):
            raise CudaLibError(
                'libminpy_cuda.so not found at %s -- build it with `python -m minpy_b200.build`; '
                'the cuda array variant has no CPU fallback' % path)
        self.path = path
        self.cdll = ctypes.CDLL(path)
        self.decls = parse_header()
        self.enums = parse_enums()
        self.cdll.mc_last_error.restype = ctypes.c_char_p
        for name, (restype, argtypes, _) in self.decls.items():
            fn = getattr(self.cdll, name)
            fn.restype = restype
            fn.argtypes = argtypes
            if restype is ctypes.c_int:
                setattr(self, name, self._checked(name, fn))
            else:
                setattr(self, name, fn)
        for k, v in self.enums.items():
            setattr(self, k, v)

    def _checked(self, name, fn):
        last_error = self.cdll.mc_last_error

        def call(*args):
            if fn(*args) != 0:
                raise CudaLibError('%s: %s' % (name, last_error().decode()))
        call.__name__ = name
        call.raw = fn
        return call


_lib = None


def get_lib() -> _Lib:
    """Load (once) and return the library; raises CudaLibError when it is unavailable."""
    global _lib
    if _lib is None:
        _lib = _Lib()
    return _lib
