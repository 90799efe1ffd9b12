"""`grad`, `grad_and_loss`, the NumPy<->MinPy converters and `Function`.

Same entry points as the reference's minpy/core.py:20-299.  `Function` wrapped an MXNet
symbol bound to an MXNet executor (core.py:148-190); here the symbol is the small graph
façade in `mxnet/symbol.py` and the executor is `minpy_b200.symbol_exec`, which lowers
Convolution / Pooling / Activation / Flatten / FullyConnected / SoftmaxOutput / BatchNorm
nodes onto the sm_100a kernels of libminpy_cuda.so.
"""
from __future__ import absolute_import
from __future__ import print_function

import functools

from minpy import array
from minpy import tape
from minpy.array_variants import ArrayType
from minpy.primitive import Primitive
from minpy.utils import log

_logger = log.get_logger(__name__)


def grad_and_loss(func, argnum=0):
    """Wrap `func` so that calling it returns `(gradients, value)`; gradients are taken with
    respect to the positional argument(s) `argnum` (an int or an iterable of ints)."""
    @functools.wraps(func)
    def wrapped(*args):
        values = tuple(array.wrap(a) for a in args)
        wanted = [argnum] if isinstance(argnum, int) else list(argnum)
        with tape.tape() as current:
            current.start_recording()
            for i in wanted:
                values[i].mark_for_bp(current)
            result = func(*values)
            current.stop_recording()
            if isinstance(result, tuple):
                target = tuple(array.wrap(r) for r in result)
            else:
                target = array.wrap(result)
            grads = current.get_gradient(tuple(values[i] for i in wanted), target)
        return (grads[0] if len(grads) == 1 else grads), result
    return wrapped


def grad(func, argnum=0):
    """Like `grad_and_loss` but returning only the gradient(s)."""
    both = grad_and_loss(func, argnum)

    @functools.wraps(func)
    def wrapped(*args):
        return both(*args)[0]
    return wrapped


class MXNetSymbolError(ValueError):
    """Malformed symbol handed to `Function`."""


class Function(object):
    """A symbol graph exposed as one differentiable primitive.

    `Function(symbol, input_shapes)(X=..., conv_weight=..., ...)`: keyword arguments are
    matched against `symbol.list_arguments()` + `list_auxiliary_states()`; parameter shapes
    come from shape inference and are reported by `get_params()`.
    """

    def __init__(self, symbol, input_shapes=None, name='mxnet_symbol'):
        for needed in ('list_arguments', 'list_outputs', 'list_auxiliary_states', 'infer_shape'):
            if not hasattr(symbol, needed):
                raise MXNetSymbolError('object %r is not a symbol (no %s)' % (symbol, needed))
        self._symbol = symbol
        self._is_train = True
        self._input_shapes = input_shapes
        self._sym_name = name
        self._plan = None
        if input_shapes is not None:
            self._infer_shape(input_shapes)

    @property
    def is_train(self):
        return self._is_train

    @is_train.setter
    def is_train(self, value):
        self._is_train = value

    def _infer_shape(self, input_shapes):
        self._input_shapes = dict(input_shapes)
        arg_shapes, out_shapes, aux_shapes = self._symbol.infer_shape(**self._input_shapes)
        names = self._symbol.list_arguments()
        self._param_shapes = dict((n, s) for n, s in zip(names, arg_shapes) if n not in input_shapes)
        self._out_shapes = dict(zip(self._symbol.list_outputs(), out_shapes))
        self._aux_shapes = dict(zip(self._symbol.list_auxiliary_states(), aux_shapes))
        self._plan = None

    def _create_prim(self, aux_holders, wanted=None):
        """Primitive running the lowered graph; its single multi-gradient VJP returns one
        gradient per argument (and zeros for auxiliary states)."""
        from minpy_b200 import symbol_exec
        if self._plan is None:
            self._plan = symbol_exec.Plan(self._symbol, self._input_shapes)
        plan = self._plan
        arg_names = self._symbol.list_arguments()
        aux_names = self._symbol.list_auxiliary_states()
        train = self._is_train

        def forward(*raw):
            run = plan.forward(raw[:len(arg_names)], raw[len(arg_names):], train, data_needs_grad=wanted)
            forward.run = run
            for holder, new in zip(aux_holders, run.new_aux):
                if holder is not None and new is not None:
                    holder[:] = array.wrap(new)      # running statistics flow back, as core.py:177-179
            outs = run.outputs
            return tuple(outs) if len(outs) > 1 else outs[0]
        forward.__name__ = self._sym_name

        def vjp_maker(ans, *raw):  # pylint: disable=unused-argument
            run = forward.run
            return lambda g: list(run.backward(g)) + [0] * len(aux_names)

        prim = Primitive(forward, ArrayType.CUDA)
        prim.def_multiple_grad(vjp_maker, tuple(range(len(arg_names) + len(aux_names))))
        return prim

    def __call__(self, **kwargs):
        if self._input_shapes is None:
            self._infer_shape(dict((k, v.shape) for k, v in kwargs.items()
                                   if k in self._symbol.list_arguments() and hasattr(v, 'shape')))
        args = [kwargs.get(n) for n in self._symbol.list_arguments()]
        aux = [kwargs.get(n) for n in self._symbol.list_auxiliary_states()]
        # arguments the tape is differentiating: the others (the image batch, labels) need no
        # gradient, so e.g. the first convolution's data-gradient is never computed
        current = tape.global_tape()
        wanted = None
        if current is not None and current.is_recording:
            wanted = set(n for n, a in zip(self._symbol.list_arguments(), args)
                         if isinstance(a, array.Value) and a.is_marked_for_bp(current))
        return self._create_prim(aux, wanted).call(args=args + aux, kwargs={})

    def get_params(self):
        return dict((name, {'shape': shape}) for name, shape in self._param_shapes.items())

    def get_output_shapes(self):
        return self._out_shapes

    def get_one_output_shape(self):
        assert len(self._out_shapes) == 1
        return list(self._out_shapes.values())[0]


def numpy_to_minpy(var):
    """NumPy array(s) -> MinPy array(s); containers are converted element-wise."""
    if isinstance(var, (tuple, list)):
        return type(var)(numpy_to_minpy(v) for v in var)
    if isinstance(var, dict):
        return dict((k, numpy_to_minpy(v)) for k, v in var.items())
    return array.wrap(var)


def minpy_to_numpy(var):
    """MinPy array(s) -> NumPy array(s); containers are converted element-wise."""
    if isinstance(var, (tuple, list)):
        return type(var)(minpy_to_numpy(v) for v in var)
    if isinstance(var, dict):
        return dict((k, minpy_to_numpy(v)) for k, v in var.items())
    return array.wrap(var).get_data(ArrayType.NUMPY)


def convert_args(func):
    """Decorator: every positional / keyword argument is wrapped into a MinPy value."""
    @functools.wraps(func)
    def wrapper(*args, **kwargs):
        return func(*[array.wrap(a) for a in args], **dict((k, array.wrap(v)) for k, v in kwargs.items()))
    return wrapper


def return_numpy(func):
    """Decorator: results are converted back to NumPy."""
    @functools.wraps(func)
    def wrapper(*args, **kwargs):
        return minpy_to_numpy(func(*args, **kwargs))
    return wrapper
