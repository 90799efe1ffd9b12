"""Gradient tape -- rewritten, same public surface as the reference's minpy/tape.py:25-271
(`Tape.start_recording / stop_recording / is_recording / add_partial_derivative /
get_gradient`, `tape()`, `global_tape()`).

What changed and why:

* Order.  The reference walks a BFS queue driven by per-array reference counts after a
  separate pruning traversal (tape.py:113-257).  Records are appended in execution order, so
  reverse recording order IS a valid reverse topological order: one backwards scan visits a
  record only after every consumer of its result has contributed, and records whose result
  never received a gradient are skipped -- which is exactly the pruning.
* Fusion.  Gradients are kept in raw form and may be lazy `ChainGrad` recipes
  (minpy_b200/lazy.py).  A run of elementwise VJPs therefore costs no launches until a
  consumer needs numbers; accumulation into an existing gradient is folded into the same
  `mc_vjp_chain` kernel.  The reference launched >= 1 array op per record plus one `add` per
  accumulation (tape.py:96-111).
* Multi-output primitives are one record whose VJP receives a tuple of gradients (None for
  outputs nothing depended on); the reference registered the same closure once per output.
"""
from __future__ import absolute_import
from __future__ import print_function

import collections
import contextlib

import numpy

from minpy import array
from minpy.array_variants import ArrayType
from minpy.utils import log
from minpy_b200.cuda_array import CudaArray
from minpy_b200.lazy import ChainGrad

_logger = log.get_logger(__name__)


def _side_join():
    from minpy_b200 import ops
    ops.side_join()


GradRecord = collections.namedtuple('GradRecord', ['grad_func', 'result', 'owner'])


def _is_zero_scalar(g):
    return isinstance(g, (int, float)) and not isinstance(g, array.Value) and g == 0


def _add(existing, new):
    """existing + new for gradients in raw form, staying on the device when either side is."""
    if isinstance(new, ChainGrad):
        if isinstance(existing, ChainGrad):
            existing = existing.materialize()
        if isinstance(existing, CudaArray) and existing.shape == new.shape:
            return new.materialize(acc=existing)
        new = new.materialize()
    if isinstance(existing, ChainGrad):
        return _add(new, existing)
    if isinstance(existing, CudaArray) or isinstance(new, CudaArray):
        from minpy_b200 import ops
        return ops.binary(ops.L().MC_ADD, existing, new)
    if isinstance(existing, array.Value) or isinstance(new, array.Value):
        return array.wrap(existing) + array.wrap(new)
    return existing + new


class Tape(object):
    """Records VJP closures during the forward pass and replays them backwards."""
    global_tape = None
    timestamp = 0

    def __init__(self):
        self._records = []
        self._consumers = collections.defaultdict(int)   # result id -> number of records reading its gradient
        self._grads = {}
        self._recording = False
        Tape.timestamp += 1

    def start_recording(self):
        self._recording = True

    def stop_recording(self):
        self._recording = False

    @property
    def is_recording(self):
        return self._recording

    def add_partial_derivative(self, grad_func, owner, result):
        """Record that `grad_func(grad of result)` yields the gradient contribution for
        `owner` (a Value, or a list aligned with a multi-gradient function; None entries are
        arguments that need no gradient).  `result` is a Value or a tuple of Values."""
        if not self._recording:
            return
        self._records.append(GradRecord(grad_func, result, owner))
        if isinstance(result, array.Value):
            self._consumers[result.id] += 1
        else:
            for res in result:
                if isinstance(res, array.Value):
                    self._consumers[res.id] += 1

    # -- backward -----------------------------------------------------------------------------
    def _seed(self, target):
        if isinstance(target, array.Number):
            self._grads[target.id] = 1.0
        elif isinstance(target, array.Array):
            if target.on_device():
                self._grads[target.id] = ChainGrad(target.shape, None, 1.0)
            else:
                self._grads[target.id] = numpy.ones(target.shape)
        elif isinstance(target, (tuple, list)):
            for sub in target:
                self._seed(sub)

    def _accumulate(self, owner, grad):
        if isinstance(owner, array.Value):
            if isinstance(grad, array.Value):
                grad = grad._get_latest_data() if isinstance(grad, array.Array) else grad.val  # pylint: disable=protected-access
            have = self._grads.get(owner.id)
            if have is not None:
                _side_join()          # an earlier contribution may still be in flight on the side stream
            self._grads[owner.id] = grad if have is None else _add(have, grad)
        elif owner is not None:
            if len(owner) != len(grad):
                _logger.fatal('Number of gradients does not match.')
            for sub_owner, sub_grad in zip(owner, grad):
                self._accumulate(sub_owner, sub_grad)

    def _grad_of(self, value):
        """Gradient stored for `value`, made concrete when several records will read it (so a
        lazy recipe is evaluated once, not once per reader)."""
        g = self._grads.get(value.id)
        if isinstance(g, ChainGrad) and self._consumers.get(value.id, 0) > 1:
            g = g.materialize()
            self._grads[value.id] = g
        return g

    def get_gradient(self, origin, target):
        """Gradients of `target` (seeded with ones) with respect to each array in `origin`;
        0.0 for origins the target does not depend on (reference tape.py:249-257)."""
        keep = set(a.id for a in origin)
        self._seed(target)
        records = self._records
        while records:
            rec = records.pop()
            result = rec.result
            if isinstance(result, array.Value):
                g = self._grad_of(result)
                if g is None:
                    continue
                ids = (result.id,)
            else:
                gs = tuple(self._grad_of(r) if isinstance(r, array.Value) else None for r in result)
                if all(x is None for x in gs):
                    continue
                g = gs
                ids = tuple(r.id for r in result if isinstance(r, array.Value))
            self._accumulate(rec.owner, rec.grad_func(g))
            # the records of one primitive call are adjacent: once the next record belongs
            # to another result, this result's gradient is dead
            nxt = records[-1].result if records else None
            if nxt is not result:
                for rid in ids:
                    if rid not in keep:
                        self._grads.pop(rid, None)
        _side_join()                  # weight gradients computed beside the chain are complete from here on
        out = []
        for arr in origin:
            g = self._grads.get(arr.id)
            if g is None:
                out.append(0.0)
                continue
            if isinstance(g, ChainGrad):
                g = g.materialize()
            out.append(array.wrap(g))
        self._grads = {}
        self._consumers.clear()
        return out


@contextlib.contextmanager
def tape():
    """Install a fresh global tape for the duration of the block."""
    Tape.global_tape = Tape()
    try:
        yield Tape.global_tape
    finally:
        Tape.global_tape = None


def global_tape():
    return Tape.global_tape
