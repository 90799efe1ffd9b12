"""Mini-batch iterators (reference: minpy/nn/io.py:14-289, itself modelled on mxnet.io).

`NDArrayIter` slices NumPy or MinPy arrays into `DataBatch`es with pad / discard / roll_over
handling of the last batch.  Additions for the B200 backend, both off unless asked for:
`num_parts` / `part_index` shard the rows across data-parallel ranks (defaults come from the
process group when one is initialised), and MinPy arrays that already live on the device are
sliced there, so a device-resident dataset costs no host->device copy per step.
"""
from __future__ import division
from __future__ import absolute_import

import pickle
from collections import OrderedDict

import numpy as np

from minpy import array


class DataBatch(object):  # pylint: disable=too-few-public-methods
    """One mini-batch: lists of data and label arrays, the number of padding rows and an index."""

    def __init__(self, data, label, pad=None, index=None):
        self.data = data
        self.label = label
        self.pad = pad
        self.index = index


class DataIter(object):
    """Iterator protocol shared by all data iterators."""

    def __init__(self):
        self.batch_size = 0

    def __iter__(self):
        return self

    def reset(self):
        """Rewind to the first batch."""

    def next(self):
        if self.iter_next():
            return DataBatch(data=self.getdata(), label=self.getlabel(), pad=self.getpad(), index=self.getindex())
        raise StopIteration

    def __next__(self):
        return self.next()

    def iter_next(self):
        """Advance; False when the epoch is over."""

    def getdata(self):
        """Data arrays of the current batch."""

    def getlabel(self):
        """Label arrays of the current batch."""

    def getindex(self):
        """Index of the current batch."""

    def getpad(self):
        """Number of padding examples in the current batch."""


_ARRAY_KINDS = (np.ndarray, array.Array)


def _init_data(data, allow_empty, default_name):
    """Canonical list of (name, array) pairs from an array, a list of arrays or a dict."""
    assert data is not None or allow_empty
    if data is None:
        data = []
    if isinstance(data, _ARRAY_KINDS):
        data = [data]
    if isinstance(data, list):
        assert allow_empty or len(data) > 0
        if len(data) == 1:
            data = OrderedDict([(default_name, data[0])])
        else:
            data = OrderedDict(('_%d_%s' % (i, default_name), d) for i, d in enumerate(data))
    if not isinstance(data, dict):
        raise TypeError('Input must be NDArray, numpy.ndarray, MinPy Array, or a list of them or dict with '
                        'them as values.')
    for key, value in data.items():
        if not isinstance(value, _ARRAY_KINDS):
            raise TypeError("Invalid type '%s' for %s, should be NDArray, numpy.ndarray, or MinPy Array." %
                            (type(value), key))
    return list(data.items())


def _default_partition():
    try:
        from minpy_b200 import parallel
        return parallel.world_size(), parallel.rank()
    except Exception:  # pylint: disable=broad-except
        return 1, 0


def _shared_order(order):
    """Rank 0's permutation on every rank of an initialised process group (else `order` unchanged)."""
    try:
        from minpy_b200 import parallel
        return parallel.broadcast_host_array(order)
    except ImportError:
        return order


class _DevicePrefetch(object):
    """Two device staging slots filled by host->device copies on the library's copy stream
    (mc_prefetch_h2d): while the step on batch i runs on the compute stream, batch i+1 is already
    travelling.  The reference slices on the host and copies inside the step (io.py:238-258 +
    array.py:388-389), which serialises PCIe and compute.  A batch handed out stays valid until two
    more batches have been drawn (its slot is then overwritten)."""

    def __init__(self):
        self.slots = [None, None]      # per slot: list of CudaArray
        self.turn = 0
        self.cursor = None             # cursor of the batch in flight
        self.pending = None
        self.keep = [None, None]       # host arrays of in-flight copies

    def forget(self):
        self.cursor, self.pending = None, None

    def start(self, cursor, host_arrays, pad):
        from minpy_b200.cuda_array import CudaArray
        from minpy_b200.ops import L
        from minpy_b200.step_graph import _dtype_on_device
        lib = L()
        slot = self.turn
        self.turn ^= 1
        hosts = [np.ascontiguousarray(h, dtype=_dtype_on_device(np.asarray(h))) for h in host_arrays]
        dev = self.slots[slot]
        if dev is None or [d.shape for d in dev] != [tuple(h.shape) for h in hosts] or \
                [d.dtype for d in dev] != [h.dtype for h in hosts]:
            dev = self.slots[slot] = [CudaArray.empty(h.shape, h.dtype) for h in hosts]
        lib.mc_prefetch_release()          # the slot's previous consumer was enqueued before this point
        for d, h in zip(dev, hosts):
            lib.mc_prefetch_h2d(d.ptr, h.ctypes.data, h.nbytes)
        self.keep[slot] = hosts            # pinned sources are read asynchronously: keep them alive
        self.cursor, self.pending = cursor, (slot, pad)

    def take(self):
        from minpy_b200.ops import L
        slot, pad = self.pending
        L().mc_prefetch_acquire()          # the compute stream waits for the copies
        self.pending = None
        return [array.wrap(d) for d in self.slots[slot]], pad


class NDArrayIter(DataIter):
    # pylint: disable=too-many-instance-attributes
    """Batches over in-memory arrays.

    data / label: an array, a list of arrays or a dict name -> array (NumPy or MinPy).
    last_batch_handle: 'pad' wraps around to fill the last batch, 'discard' drops it,
    'roll_over' carries the remainder into the next epoch.
    """

    def __init__(self, data, label=None, batch_size=1, shuffle=False, last_batch_handle='pad',
                 num_parts=None, part_index=None, device_prefetch=False):
        # pylint: disable=too-many-arguments
        super(NDArrayIter, self).__init__()
        # not in the reference: copy batch i+1 to the device on a second stream while step i computes
        # (see _DevicePrefetch); the batches handed out are then device arrays
        self._prefetch = _DevicePrefetch() if device_prefetch else None
        self.data = _init_data(data, allow_empty=False, default_name='data')
        self.label = _init_data(label, allow_empty=True, default_name='softmax_label')

        if num_parts is None:
            num_parts, part_index = _default_partition()
        if shuffle:   # once, at construction, with the global NumPy RNG (reference :173-177)
            order = np.arange(self.data[0][1].shape[0])
            np.random.shuffle(order)
            if num_parts > 1:
                # data parallel: the strided shards below partition the data only if every rank
                # permutes with the SAME order -- rank 0's (each process has its own unseeded RNG)
                order = _shared_order(order)
            self.data = [(k, v[order]) for k, v in self.data]
            self.label = [(k, v[order]) for k, v in self.label]

        if num_parts > 1:   # data-parallel shard: rows part_index, part_index + num_parts, ...
            self.data = [(k, v[part_index::num_parts]) for k, v in self.data]
            self.label = [(k, v[part_index::num_parts]) for k, v in self.label]
        self.num_parts, self.part_index = num_parts, part_index or 0

        self.data_list = [v for _, v in self.data] + [v for _, v in self.label]
        self.num_source = len(self.data_list)
        self.num_iterations_per_epoch = self.data[0][1].shape[0] / batch_size   # a float, as the reference

        if last_batch_handle == 'discard':
            keep = self.data_list[0].shape[0] - self.data_list[0].shape[0] % batch_size
            self.data = [(k, v[:keep]) for k, v in self.data]
            self.label = [(k, v[:keep]) for k, v in self.label]
        self.num_data = self.data_list[0].shape[0]
        assert self.num_data >= batch_size, 'batch_size need to be smaller than data size.'
        self.cursor = -batch_size
        self.batch_size = batch_size
        self.last_batch_handle = last_batch_handle

    @property
    def provide_data(self):
        return [(k, tuple([self.batch_size] + list(v.shape[1:]))) for k, v in self.data]

    @property
    def provide_label(self):
        return [(k, tuple([self.batch_size] + list(v.shape[1:]))) for k, v in self.label]

    def hard_reset(self):
        """Rewind, forgetting any rolled-over remainder."""
        self.cursor = -self.batch_size

    def reset(self):
        if self.last_batch_handle == 'roll_over' and self.cursor > self.num_data:
            self.cursor = -self.batch_size + (self.cursor % self.num_data) % self.batch_size
        else:
            self.cursor = -self.batch_size

    def iter_next(self):
        self.cursor += self.batch_size
        return self.cursor < self.num_data

    def next(self):
        if self._prefetch is not None:
            return self._next_prefetched()
        if self.iter_next():
            return DataBatch(data=self.getdata(), label=self.getlabel(), pad=self.getpad(), index=None)
        raise StopIteration

    def _host_batch(self, cursor):
        """Host arrays (data + label tensors) of the batch starting at `cursor`, and its pad: VIEWS of
        the sources (a pinned source is then copied to the device without touching the host again)."""
        lo, hi = cursor, cursor + self.batch_size
        out = []
        for _, v in self.data + self.label:
            if isinstance(v, array.Array):
                v = v.get_data(array.ArrayType.NUMPY)
            if hi <= self.num_data:
                out.append(v[lo:hi])
            else:
                out.append(np.concatenate((v[lo:], v[:hi - self.num_data]), axis=0))
        pad = hi - self.num_data if (self.last_batch_handle == 'pad' and hi > self.num_data) else 0
        return out, pad

    def _next_prefetched(self):
        pf = self._prefetch
        if not self.iter_next():
            if pf.cursor != 0:                             # keep a first batch already travelling (below)
                pf.forget()
            raise StopIteration
        if pf.cursor != self.cursor:                       # first batch ever (or after an unusual reset)
            pf.start(self.cursor, *self._host_batch(self.cursor))
        arrays, pad = pf.take()
        nxt = self.cursor + self.batch_size
        if nxt < self.num_data:
            pf.start(nxt, *self._host_batch(nxt))
        elif self.last_batch_handle != 'roll_over':
            # last batch of the epoch: the order is fixed at construction (shuffle happens once, :173-177),
            # so the next epoch's first batch is known -- send it now instead of stalling after reset()
            pf.start(0, *self._host_batch(0))
        nd = len(self.data)
        return DataBatch(data=arrays[:nd], label=arrays[nd:], pad=pad, index=None)

    def _getdata(self, source):
        assert self.cursor < self.num_data, 'DataIter needs reset.'
        lo, hi = self.cursor, self.cursor + self.batch_size
        if hi <= self.num_data:
            return [v[lo:hi] if isinstance(v, array.Array) else array.wrap(v[lo:hi]) for _, v in source]
        wrap_around = hi - self.num_data
        out = []
        for _, v in source:
            if isinstance(v, array.Array):
                v = v.asnumpy()
            out.append(array.wrap(np.concatenate((v[lo:], v[:wrap_around]), axis=0)))
        return out

    def getdata(self):
        return self._getdata(self.data)

    def getlabel(self):
        return self._getdata(self.label)

    def getpad(self):
        if self.last_batch_handle == 'pad' and self.cursor + self.batch_size > self.num_data:
            return self.cursor + self.batch_size - self.num_data
        return 0

    def getindex(self):
        return self.cursor / self.batch_size

    def getsubiter(self, num_samples):
        """A shuffled iterator over a random subset of `num_samples` rows."""
        order = np.arange(self.data[0][1].shape[0])
        np.random.shuffle(order)
        chosen = order[0:num_samples]
        return NDArrayIter([v[chosen] for _, v in self.data], [v[chosen] for _, v in self.label],
                           self.batch_size, True, num_parts=1, part_index=0)

    def getnumiterations(self):
        return self.num_iterations_per_epoch


def save_data_labels(data_vec, label_vec, file_name):
    """Pickle {'data': ..., 'labels': ...} to `file_name`."""
    with open(file_name, 'wb') as f:
        pickle.dump({'data': data_vec, 'labels': label_vec}, f, protocol=pickle.HIGHEST_PROTOCOL)


def load_data_labels(file_name):
    """Inverse of `save_data_labels`: returns (data, labels)."""
    with open(file_name, 'rb') as f:
        blob = pickle.load(f)
    return blob['data'], blob['labels']
