"""`ModelBase`: parameter bookkeeping a `Solver` understands (reference:
minpy/nn/model.py:14-163).  Checkpoints use HDF5 when h5py is importable and NumPy's .npz
container otherwise (same keys: `param_<name>`, `aux_param_<name>`)."""
import numpy

# pylint: disable=invalid-name


class ParamsNameNotFoundError(ValueError):
    """A checkpoint lacks a parameter the model declares."""


class UnknownAccessModeError(ValueError):
    """Kept for API compatibility."""


class ModelBase(object):
    """Subclasses declare parameters with `add_param` and implement `forward` and `loss`."""

    def __init__(self):
        self.params = {}
        self.param_configs = {}
        self.aux_params = {}
        self.aux_param_configs = {}

    def add_param(self, name, shape, **kwargs):
        """Declare a trainable parameter; extra keywords (init_rule, init_config, ...) are kept in
        its config.  Returns self for chaining."""
        assert name not in self.param_configs, 'Duplicate parameter name %s' % name
        config = {'shape': shape}
        config.update(kwargs)
        self.param_configs[name] = config
        return self

    def add_params(self, param_dict):
        """Declare several parameters from {name: config}."""
        for name, config in param_dict.items():
            assert name not in self.param_configs, 'Duplicate parameter name %s' % name
            self.param_configs[name] = config
        return self

    def add_aux_param(self, name, value):
        """Declare a non-trained state tensor (e.g. batchnorm running statistics)."""
        assert name not in self.aux_param_configs, 'Duplicate auxiliary parameter name %s' % name
        self.aux_param_configs[name] = value
        return self

    def forward_batch(self, batch, mode):
        """Forward on a DataBatch; by default only its first data array is used."""
        return self.forward(batch.data[0], mode)

    def loss_batch(self, batch, forward_outputs):
        """Loss on a DataBatch; by default only its first label array is used."""
        return self.loss(forward_outputs, batch.label[0])

    def forward(self, X, mode):
        raise NotImplementedError()

    def loss(self, predict, y):
        raise NotImplementedError()

    # -- checkpoints ----------------------------------------------------------------------
    def _named_tensors(self):
        items = [('param_%s' % k, v) for k, v in self.params.items()]
        items += [('aux_param_%s' % k, v) for k, v in self.aux_params.items()]
        return items

    def save(self, prefix):
        """Write `<prefix>.params`."""
        path = '%s.params' % prefix
        # auxiliary states that do not exist yet (batchnorm running statistics before the first step) are skipped
        tensors = dict((k, numpy.asarray(v.asnumpy() if hasattr(v, 'asnumpy') else v))
                       for k, v in self._named_tensors() if v is not None)
        try:
            import h5py
        except ImportError:
            with open(path, 'wb') as f:
                numpy.savez(f, **tensors)
            return
        with h5py.File(path, 'w') as hf:
            for key, value in tensors.items():
                hf.create_dataset(key, data=value)

    def load(self, prefix):
        """Read `<prefix>.params` into the existing parameter arrays (in place)."""
        path = '%s.params' % prefix
        try:
            import h5py
            with h5py.File(path, 'r') as hf:
                stored = dict((k, numpy.array(hf.get(k))) for k in hf.keys())
        except (ImportError, OSError):
            with numpy.load(path) as z:
                stored = dict((k, z[k]) for k in z.files)
        for key, tensor in self._named_tensors():
            if tensor is None:
                if key in stored:          # an auxiliary state the checkpoint has and this model has not created yet
                    import minpy.numpy as mnp
                    self.aux_params[key[len('aux_param_'):]] = mnp.array(stored[key])
                continue
            if key not in stored:
                raise ParamsNameNotFoundError('%s not found in %s' % (key, path))
            tensor[:] = stored[key]
