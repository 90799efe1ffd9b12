"""Device contexts: `Context('cpu'|'gpu', id)`, `with` scoping and a process-wide default.

Same surface as the reference's minpy/context.py:17-153, without the `mxnet.Context` it
wrapped (context.py:55-56): for the `cuda` array variant a gpu context names the CUDA device
the primitive runs on; a cpu context leaves the device choice to LOCAL_RANK (one process
per GPU), because the cuda variant has no host execution path.
"""
from __future__ import absolute_import

_TYPE_NAMES = {1: 'cpu', 2: 'gpu', 3: 'cpu_pinned'}
_TYPE_IDS = dict((v, k) for k, v in _TYPE_NAMES.items())


class Context(object):
    """A (device type, device id) pair; usable as a `with` block to scope the default."""
    default_ctx = None
    devtype2str = _TYPE_NAMES
    devstr2type = _TYPE_IDS

    def __init__(self, device_type, device_id=0):
        if isinstance(device_type, Context):
            device_type, device_id = device_type.device_type, device_type.device_id
        if device_type not in _TYPE_IDS:
            raise ValueError('unknown device type %r (expected one of %s)' % (device_type, sorted(_TYPE_IDS)))
        self.device_typeid = _TYPE_IDS[device_type]
        self.device_id = int(device_id)
        self._saved = []

    @property
    def device_type(self):
        return _TYPE_NAMES[self.device_typeid]

    def cuda_device(self):
        """CUDA ordinal for this context, or None when the context does not pin one."""
        return self.device_id if self.device_typeid == 2 else None

    def __eq__(self, other):
        return (isinstance(other, Context) and other.device_typeid == self.device_typeid
                and other.device_id == self.device_id)

    def __ne__(self, other):
        return not self.__eq__(other)

    def __hash__(self):
        return hash((self.device_typeid, self.device_id))

    def __str__(self):
        return '%s(%d)' % (self.device_type, self.device_id)

    __repr__ = __str__

    def __enter__(self):
        self._saved.append(Context.default_ctx)
        Context.default_ctx = self
        return self

    def __exit__(self, exc_type, exc, tb):
        Context.default_ctx = self._saved.pop()


Context.default_ctx = Context('cpu', 0)


def cpu(device_id=0):
    """Shorthand for Context('cpu', device_id)."""
    return Context('cpu', device_id)


def gpu(device_id=0):
    """Shorthand for Context('gpu', device_id)."""
    return Context('gpu', device_id)


def current_context():
    return Context.default_ctx


def set_context(ctx):
    Context.default_ctx = ctx
