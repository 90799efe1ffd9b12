"""`Value`, `Number`, `Array` and `wrap`: the objects user code holds.

Same contract as the reference's minpy/array.py:26-534 -- operators redirect to the mocked
`minpy.numpy` namespace so that they are dispatched (and taped) like any other primitive;
an `Array` keeps one raw array per variant plus a marker telling which copy is current --
generalised from the reference's hard-wired NumPy<->MXNet pair (array.py:375-390) to any
number of variants with NumPy as the hub, and with indexing kept on the device when the
current copy lives there (the reference always indexed on the host, array.py:432-443).
"""
from __future__ import absolute_import
from __future__ import print_function

import collections.abc
import itertools

import numpy

from minpy.array_variants import ArrayType, array_types, number_types
from minpy.context import current_context
from minpy.utils import log
from minpy_b200.cuda_array import CudaArray

_logger = log.get_logger(__name__)


class Value(object):
    """Anything that can flow through primitives and be differentiated."""
    _ns = None                      # the `minpy.numpy` namespace, set when it is built
    _ids = itertools.count(0)

    def __init__(self, context=None):
        self._bp_timestamp = -1
        self._minpy_value_id = next(Value._ids)
        self._context = current_context() if context is None else context

    # -- autograd marking -----------------------------------------------------------------
    def is_marked_for_bp(self, tape):
        return tape is not None and tape.is_recording and self._bp_timestamp == tape.timestamp

    def mark_for_bp(self, tape):
        self._bp_timestamp = tape.timestamp

    def wait_to_read(self):
        """Block until the value is computed (no-op for host values)."""

    @property
    def context(self):
        return self._context

    @property
    def id(self):  # pylint: disable=invalid-name
        return self._minpy_value_id

    def __hash__(self):
        return id(self)


def _install_operators():
    """Route Python operators to namespace functions.  In-place forms are NOT in place: they
    rebind to a fresh result (reference array.py:199-221), which is what lets optimizer code
    such as `w -= lr * dw` stay differentiable-safe."""
    binary = {'add': 'add', 'sub': 'subtract', 'mul': 'multiply', 'div': 'divide',
              'truediv': 'true_divide', 'mod': 'mod', 'pow': 'power'}
    compare = {'eq': 'equal', 'ne': 'not_equal', 'lt': 'less', 'gt': 'greater',
               'le': 'less_equal', 'ge': 'greater_equal'}
    unary = {'neg': 'negative', 'abs': 'abs'}
    unsupported = ['pos', 'invert', 'floor', 'ceil', 'trunc', 'floordiv', 'divmod', 'lshift',
                   'rshift', 'and', 'or', 'xor', 'rfloordiv', 'rlshift', 'rrshift', 'rand', 'ror',
                   'rxor', 'ifloordiv', 'ilshift', 'irshift', 'iand', 'ior', 'ixor', 'cmp']

    def forward(fname):
        return lambda self, other: getattr(Value._ns, fname)(self, other)

    def reflected(fname):
        return lambda self, other: getattr(Value._ns, fname)(other, self)

    for op, fname in binary.items():
        setattr(Value, '__%s__' % op, forward(fname))
        setattr(Value, '__r%s__' % op, reflected(fname))
        setattr(Value, '__i%s__' % op, forward(fname))
    for op, fname in compare.items():
        setattr(Value, '__%s__' % op, forward(fname))
    for op, fname in unary.items():
        setattr(Value, '__%s__' % op, (lambda f: lambda self: getattr(Value._ns, f)(self))(fname))
    Value.__rdivmod__ = reflected('mod')

    def refuse(name):
        def method(self, *args):
            raise NotImplementedError('operator __%s__ is not supported on MinPy values' % name)
        return method
    for op in unsupported:
        setattr(Value, '__%s__' % op, refuse(op))
    Value.__round__ = refuse('round')


_install_operators()


class Number(Value, float):
    """A Python float that can sit on the tape."""
    __slots__ = ['_val']

    def __new__(cls, val, context=None):  # pylint: disable=unused-argument
        return float.__new__(cls, val)

    def __init__(self, val, context=None):
        Value.__init__(self, context)
        self._val = val

    def __str__(self):
        return str(self._val)

    def __repr__(self):
        return repr(self._val)

    __hash__ = Value.__hash__

    def get_data(self, dtype):  # pylint: disable=unused-argument
        return self._val

    get_data_mutable = get_data

    def asnumpy(self):
        return self._val

    @property
    def val(self):
        return self._val


class Array(Value):
    """N-d array with one raw copy per variant and lazy synchronisation between them."""
    __slots__ = ['_data', '_latest_version', '_dtype']
    __array_priority__ = 100.0

    def __init__(self, data, atype, context=None):
        Value.__init__(self, context)
        self._data = {atype: data}
        self._latest_version = atype
        self._dtype = data.dtype

    # -- raw data access ------------------------------------------------------------------
    def has_type(self, atype):
        return atype in self._data

    def _synchronize_data(self, want):
        """Bring the `want` copy up to date from the current one (host<->device copy)."""
        src = self._latest_version
        if src == ArrayType.CUDA and want == ArrayType.NUMPY:
            _logger.info('device->host copy of Array %d, shape %s', self.id, self.shape)
            self._data[ArrayType.NUMPY] = self._data[ArrayType.CUDA].asnumpy()
        elif src == ArrayType.NUMPY and want == ArrayType.CUDA:
            _logger.info('host->device copy of Array %d, shape %s', self.id, self.shape)
            device = self._context.cuda_device()
            if device is not None:
                from minpy_b200.cuda_array import ensure_device
                ensure_device(device)
            self._data[ArrayType.CUDA] = CudaArray.from_numpy(self._data[ArrayType.NUMPY])
        else:
            raise NotImplementedError('no conversion from variant %r to %r' % (src, want))
        # every copy present is now current
        self._latest_version = None

    def get_data(self, dtype):
        latest = self._latest_version
        if latest is not None and latest != dtype:
            self._synchronize_data(dtype)
        elif latest is None and dtype not in self._data:
            self._latest_version = next(iter(self._data))
            self._synchronize_data(dtype)
        return self._data[dtype]

    def get_data_mutable(self, dtype):
        data = self.get_data(dtype)
        self._latest_version = dtype
        for other in [k for k in self._data if k != dtype]:
            del self._data[other]       # stale after the mutation
        return data

    def _get_latest_data(self):
        if self._latest_version is not None:
            return self._data[self._latest_version]
        for atype in (ArrayType.NUMPY, ArrayType.CUDA):
            if atype in self._data:
                return self._data[atype]
        raise RuntimeError('Array holds no data')

    def on_device(self):
        """True when the current copy is a device array (no host copy would be fresher)."""
        return ArrayType.CUDA in self._data and self._latest_version in (None, ArrayType.CUDA)

    # -- ndarray-like surface -------------------------------------------------------------
    def __str__(self):
        return str(self.get_data(ArrayType.NUMPY))

    def __repr__(self):
        return repr(self.get_data(ArrayType.NUMPY))

    def __len__(self):
        return self.shape[0]

    def __float__(self):
        return float(self.get_data(ArrayType.NUMPY).reshape(-1)[0]) if self.size == 1 else float(
            self.get_data(ArrayType.NUMPY))

    def __bool__(self):
        return bool(self.get_data(ArrayType.NUMPY))

    @property
    def dtype(self):
        return self._dtype

    @property
    def shape(self):
        return self._get_latest_data().shape

    @property
    def size(self):
        return self._get_latest_data().size

    @property
    def ndim(self):
        return len(self.shape)

    @property
    def T(self):  # pylint: disable=invalid-name
        return Value._ns.transpose(self)

    def asnumpy(self):
        """A host copy as numpy.ndarray (synchronises the device)."""
        return numpy.array(self.get_data(ArrayType.NUMPY))

    def wait_to_read(self):
        if self.has_type(ArrayType.CUDA):
            self._data[ArrayType.CUDA].wait_to_read()

    def reshape(self, *args, **kwargs):
        """`a.reshape(2, 3)` or `a.reshape((2, 3))`; C order only."""
        if kwargs.get('order', 'C') != 'C':
            raise NotImplementedError('Orders other than C are not currently supported.')
        if len(args) == 1 and isinstance(args[0], collections.abc.Iterable):
            shape = tuple(args[0])
        else:
            shape = tuple(args)
        return Value._ns.reshape(self, shape)

    def dot(self, other, out=None):
        if out is not None:
            raise ValueError('Out option is not supported.')
        return Value._ns.dot(self, other)

    def argmax(self, axis=None, out=None):
        if out is not None:
            raise ValueError('Out option is not supported.')
        return Value._ns.argmax(self, axis)

    # -- indexing -------------------------------------------------------------------------
    def __getitem__(self, index):
        """Differentiable indexing through the `_minpy_getitem` primitive."""
        return Value._ns._minpy_getitem(self, _normalize_index(index))  # pylint: disable=protected-access

    def __setitem__(self, index, val):
        """In-place write on the host copy; breaks the gradient chain (reference
        array.py:445-458)."""
        host_val = val
        if isinstance(val, Value):
            host_val = val.get_data(ArrayType.NUMPY)
        elif isinstance(val, CudaArray):
            host_val = val.asnumpy()
        host = self.get_data_mutable(ArrayType.NUMPY)
        host[_host_index(index)] = host_val

    def __delitem__(self, index):
        raise NotImplementedError('arrays do not support item deletion')

    def __iter__(self):
        for i in range(len(self)):
            yield self[i]


def _index_item(item, to_host):
    if isinstance(item, (slice, type(Ellipsis), type(None), int)):
        return item
    if isinstance(item, Array):
        if not to_host and item.on_device():
            return item.get_data(ArrayType.CUDA)       # labels already on the device stay there
        item = item.get_data(ArrayType.NUMPY)
    elif isinstance(item, Number):
        item = item.val
    elif isinstance(item, CudaArray):
        if not to_host:
            return item
        item = item.asnumpy()
    if isinstance(item, numpy.ndarray) and item.dtype.kind != 'b':
        return item.astype(numpy.int32)                # reference array.py:494-495
    if isinstance(item, (list, tuple)):
        return numpy.asarray(item)
    if isinstance(item, (float, numpy.floating)):
        return int(item)
    return item


def _normalize_index(index):
    """Index as the `_minpy_getitem` primitives expect it."""
    if isinstance(index, tuple):
        return tuple(_index_item(i, False) for i in index)
    return _index_item(index, False)


def _host_index(index):
    if isinstance(index, tuple):
        return tuple(_index_item(i, True) for i in index)
    return _index_item(index, True)


_NUMBER_TYPES = tuple(set(t for tys in number_types.values() for t in tys))


def wrap(data):
    """Wrap raw data in its MinPy class: arrays of any variant -> `Array`, numbers ->
    `Number`; `Value`s, None and everything else pass through (reference array.py:514-534)."""
    if data is None or isinstance(data, Value):
        return data
    if isinstance(data, numpy.ndarray):
        return Array(data, ArrayType.NUMPY)
    if isinstance(data, CudaArray):
        return Array(data, ArrayType.CUDA)
    if type(data) in _NUMBER_TYPES:  # pylint: disable=unidiomatic-typecheck
        return Number(data)
    return data
