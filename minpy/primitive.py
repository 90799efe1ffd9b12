"""`Primitive`: one forward function of one array variant, plus its VJP makers.

Keeps the protocol of the reference's minpy/primitive.py:65-483 -- `Primitive(func, type,
mutate_args, mutate_kw, type_str)`, `def_grad / def_grad_kw / def_grad_zero /
def_multiple_grad`, `gradable`, `call(args, kwargs)`, `customop` -- so variant modules written
for the reference register unchanged.  What differs is underneath: raw conversion knows N
variants, multi-output primitives are taped as ONE record receiving a tuple of gradients, and
the closure handed to the tape accepts whatever form the rewritten tape stores a gradient in
(device array, host array, Python scalar or a lazy `ChainGrad` recipe).
"""
from __future__ import absolute_import
from __future__ import print_function

import collections

import numpy

from minpy import array
from minpy import tape as tape_mod
from minpy.array import Value
from minpy.array_variants import ArrayType, variants, variants_repr
from minpy.utils import log
from minpy.context import Context
from minpy_b200.cuda_array import CudaArray, _state as _cuda_state, enter_context_device, ensure_device
from minpy_b200.lazy import ChainGrad

_logger = log.get_logger(__name__)

GradFunc = collections.namedtuple('GradFunc', ['f', 'multi_grad_indices'])


class NoGradientFuncError(ValueError):
    """Backward reached an argument whose partial derivative was never defined."""

    def __init__(self, func_name, arg):
        what = 'keyword argument "%s"' % arg if isinstance(arg, str) else 'argument %s' % (arg,)
        super(NoGradientFuncError, self).__init__(
            'Partial derivative of func "%s" on %s is not defined.' % (func_name, what))
        self.func_name = func_name
        self.arg = arg


class FakeGradFunc(object):
    """Stand-in recorded for an undefined partial derivative: raising is deferred to the
    backward pass, and only if that path is actually needed (reference :23-62)."""

    def __init__(self, func_name, arg):
        self.func_name = func_name
        self.arg = arg

    def __call__(self, grad_value):
        raise NoGradientFuncError(self.func_name, self.arg)


def to_raw(value, vtype):
    """Gradient / value in any stored form -> raw data of variant `vtype`."""
    if isinstance(value, tuple):
        return tuple(to_raw(v, vtype) for v in value)
    if isinstance(value, Value):
        return value.get_data(vtype)
    if vtype == ArrayType.CUDA:
        if isinstance(value, numpy.ndarray) and value.ndim > 0:
            return CudaArray.from_numpy(value)
        return value
    if isinstance(value, ChainGrad):
        value = value.materialize()
    if isinstance(value, CudaArray):
        return value.asnumpy()
    return value


class _TapedGrad(object):
    """What the tape calls: converts the stored gradient to this primitive's raw type."""
    __slots__ = ('fn', 'vtype', 'dev')

    def __init__(self, fn, vtype):
        self.fn = fn
        self.vtype = vtype
        # the VJP runs where the forward ran, also when backward is called outside the `with gpu(i):` scope
        self.dev = _cuda_state['device'] if vtype == ArrayType.CUDA else None

    def __call__(self, grad_value):
        if self.dev is not None and self.dev != _cuda_state['device']:
            ensure_device(self.dev)
        return self.fn(to_raw(grad_value, self.vtype))


class Primitive(object):
    """Forward function + gradient definitions for one array variant."""
    __slots__ = ['_func', '_grad_func', '_grad_func_kw', '_type', '_mutate_args', '_mutate_kw',
                 '_type_str', '__name__']

    def __init__(self, func, ty, mutate_args=None, mutate_kw=None, type_str=None):
        self._func = func
        self._grad_func = {}
        self._grad_func_kw = {}
        self._type = ty
        self._mutate_args = frozenset(mutate_args or ())
        self._mutate_kw = frozenset(mutate_kw or ())
        self._type_str = type_str if type_str is not None else variants_repr[ty] + ' op'
        self.__name__ = getattr(func, '__name__', type(func).__name__)

    @property
    def type(self):
        return self._type

    @property
    def typestr(self):
        return variants_repr[self._type]

    def __str__(self):
        return self.__name__

    def __call__(self, *args, **kwargs):
        return self.call(args, kwargs)

    # -- argument conversion ----------------------------------------------------------------
    def _raw(self, arg, mutate):
        if isinstance(arg, Value):
            raw = arg.get_data_mutable(self._type) if mutate else arg.get_data(self._type)
            if self._type == ArrayType.CUDA and isinstance(raw, CudaArray) and raw.device != _cuda_state['device']:
                # an operand that lives on another device than the `with gpu(i):` scope: a transient copy
                # (MXNet raised here; the array itself stays where it is, minpy/context.py:92-98)
                if mutate:
                    raise ValueError('cannot mutate an array that lives on gpu(%d) from a gpu(%d) scope' % (
                        raw.device, _cuda_state['device']))
                raw = raw.to_device(_cuda_state['device'])
            return raw
        if isinstance(arg, list) and arg and isinstance(arg[0], Value):
            return [self._raw(item, mutate) for item in arg]     # np.stack / np.concatenate
        if self._type == ArrayType.CUDA:
            if isinstance(arg, numpy.ndarray) and arg.ndim > 0 and arg.dtype.kind in 'fiub':
                return CudaArray.from_numpy(arg)
        elif isinstance(arg, CudaArray):
            return arg.asnumpy()
        return arg

    def _convert_args(self, args, kwargs):
        margs, mkw = self._mutate_args, self._mutate_kw
        arg_values = tuple(self._raw(a, i in margs) for i, a in enumerate(args))
        kwarg_values = dict((k, self._raw(a, k in mkw)) for k, a in kwargs.items()) if kwargs else {}
        return arg_values, kwarg_values

    # -- the call -----------------------------------------------------------------------------
    def call(self, args, kwargs):
        """Run the forward function on raw data, wrap the result and, while a tape records,
        register the VJP closures of every argument that is being differentiated."""
        if self._type == ArrayType.CUDA:
            ctx = Context.default_ctx
            if ctx.device_typeid == 2 or _cuda_state['device'] != _cuda_state['home']:
                enter_context_device(ctx)     # `with gpu(i):` places the primitive on device i
        arg_values, kwarg_values = self._convert_args(args, kwargs)
        result_value = self._func(*arg_values, **kwarg_values)
        if isinstance(result_value, tuple):
            result = tuple(array.wrap(r) for r in result_value)
        else:
            result = array.wrap(result_value)

        current = tape_mod.global_tape()
        if current is None or not current.is_recording:
            return result
        bp_idx = [i for i, a in enumerate(args) if isinstance(a, Value) and a.is_marked_for_bp(current)]
        bp_kw = [k for k, a in kwargs.items() if isinstance(a, Value) and a.is_marked_for_bp(current)]
        if not bp_idx and not bp_kw:
            return result

        results = result if isinstance(result, tuple) else (result,)
        for res in results:
            if isinstance(res, Value):
                res.mark_for_bp(current)

        done = set()
        for i in bp_idx:
            if i in done:
                continue
            done.add(i)
            rec = self._grad_func.get(i)
            if rec is None:
                grad_func, owner = FakeGradFunc(self._type_str + ' ' + self.__name__, i), args[i]
            else:
                grad_func = _TapedGrad(rec.f(result_value, *arg_values, **kwarg_values), self._type)
                if rec.multi_grad_indices is None:
                    owner = args[i]
                else:
                    owner = []
                    for j in rec.multi_grad_indices:
                        done.add(j)
                        a = args[j] if j < len(args) else None
                        owner.append(a if isinstance(a, Value) and a.is_marked_for_bp(current) else None)
            current.add_partial_derivative(grad_func, owner, result)
        for k in bp_kw:
            rec = self._grad_func_kw.get(k)
            if rec is None:
                grad_func = FakeGradFunc(self._type_str + ' ' + self.__name__, k)
            else:
                grad_func = _TapedGrad(rec.f(result_value, *arg_values, **kwarg_values), self._type)
            # the reference reuses a stale `owner` here (primitive.py:350-375);
            # the keyword argument itself is the owner
            current.add_partial_derivative(grad_func, kwargs[k], result)
        return result

    # -- gradient definitions ---------------------------------------------------------------
    def def_grad(self, func, argnum=0):
        """`func(ans, *args, **kwargs)` returns the VJP `g -> gradient` of argument `argnum`."""
        self._grad_func[argnum] = GradFunc(f=func, multi_grad_indices=None)
        return self

    def def_grad_kw(self, func, key):
        self._grad_func_kw[key] = GradFunc(f=func, multi_grad_indices=None)
        return self

    def def_grad_zero(self, argnum=0):
        return self.def_grad(lambda *args, **kwargs: lambda g: 0.0, argnum)

    def def_multiple_grad(self, func, argnums):
        """One VJP returning a list of gradients aligned with `argnums`."""
        assert isinstance(argnums, tuple), 'Indexes must be of type tuple.'
        assert len(argnums) == len(set(argnums)), 'Duplicate entries.'
        for argnum in argnums:
            assert isinstance(argnum, int), 'Indexes must be tuple of integers.'
            self._grad_func[argnum] = GradFunc(f=func, multi_grad_indices=argnums)
        return self

    def gradable(self, bp_args, bp_kwargs):
        return (all(i in self._grad_func for i in bp_args) and
                all(k in self._grad_func_kw for k in bp_kwargs))


def customop(op_type):
    """Decorator turning a raw function into a user-defined primitive of the given variant
    ('numpy' or 'cuda'; 'mxnet' is accepted as an alias of 'cuda')."""
    op_type = {'mxnet': 'cuda'}.get(op_type, op_type)
    if op_type not in variants:
        raise ValueError('Wrong type for CustomOp. Must be either "numpy" or "cuda"')

    def decorator(func):
        return Primitive(func, variants[op_type], type_str='customized op')
    return decorator
