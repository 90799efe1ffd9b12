"""Object-oriented model building: `Module` / `Sequential` / `Layer` / `Model` / `Updater`.

Same user-facing surface as the reference's minpy/nn/model_builder.py:39-744 (the second entry point
beside `ModelBase` + `Solver`; used by examples/nn/resnet.py, vgg.py, rnn_on_mnist.py):

    class Net(Model):
        def __init__(self):
            super(Net, self).__init__(loss='softmax_loss')
            self.body = Sequential(FullyConnected(num_hidden=64), ReLU(), FullyConnected(num_hidden=10))
        def forward(self, data, mode='training'):
            return self.body(data)

    net = Net()
    updater = Updater(net, update_rule='sgd_momentum', learning_rate=0.1)
    loss = net(data, labels=labels)        # forward under a fresh tape
    updater(net.backward())                # {parameter name: gradient} -> update rules of nn/optim.py

* Modules assigned as attributes of a `Model` (directly, inside tuples / lists, or as operands of
  `a + b` / `Sequential`) are registered with it; a `Layer` creates its parameters in
  `model.params` the first time it is called, with shapes taken from its inputs
  (`param_shapes(*input_shapes)`) and the init rule of its per-parameter `init_configs`.
* `Model.__call__` records the forward (+ loss) on a tape of its own with every parameter marked;
  `backward()` returns the gradients by parameter name; `Updater(grad_dict)` applies, per parameter, the
  rule named in its `update_configs` (`update_rule`, `learning_rate`, ...; state lives in the config).

Differences from the reference, all deliberate: a layer really is initialised once (the reference sets
`_to_initialized` and so re-enters the initialisation branch on every call, model_builder.py:282-299),
`Sequential` is a real sequence (`len`, indexing, iteration work), `Model.grad_and_loss` is implemented.
"""
from __future__ import absolute_import
from __future__ import print_function

import collections
import copy
import functools
import operator

import numpy

import minpy.array
import minpy.nn.init
import minpy.nn.layers
import minpy.nn.model
import minpy.nn.optim
import minpy.tape

_counters = collections.Counter()


def _auto_name(kind):
    index = _counters[kind]
    _counters[kind] += 1
    return '%s%d' % (kind, index)


def _is_array(value):
    return isinstance(value, (minpy.array.Array, numpy.ndarray))


def _is_minpy_value(value):
    return isinstance(value, (minpy.array.Array, minpy.array.Number))


# ------------------------------------------------------------------------------ modules ----
class Module(object):
    """Anything callable on arrays that may own parameters of a model."""
    _module_name = None

    def __init__(self, name=None):
        if self._module_name is None:
            raise NotImplementedError('%s does not define _module_name' % type(self).__name__)
        self._name = _auto_name(self._module_name) if name is None else name

    @property
    def name(self):
        return self._name

    def forward(self, *args, **kwargs):
        raise NotImplementedError()

    def __call__(self, *args, **kwargs):
        raise NotImplementedError()

    def __repr__(self):
        return str(self)

    def __str__(self):
        return self._name

    def __add__(self, other):
        return Add(self, other)

    def __sub__(self, other):
        return Sub(self, other)

    def __mul__(self, other):
        return Mul(self, other)

    def training(self):
        pass

    def inference(self):
        pass

    def _affiliate_to(self, model):
        """Bind to `model`: parameters live in model.params, update configs in model._update_configs;
        containers pass the call on to what they contain."""
        raise NotImplementedError()


class Container(Module):
    """A module made of modules; owns no parameters itself."""

    def __call__(self, *args, **kwargs):
        return self.forward(*args, **kwargs)

    def _children(self):
        return ()

    def training(self):
        for child in self._children():
            child.training()

    def inference(self):
        for child in self._children():
            child.inference()

    def _affiliate_to(self, model):
        for child in self._children():
            child._affiliate_to(model)  # pylint: disable=protected-access

    def param_shapes(self, input_shape):  # pylint: disable=unused-argument
        return {}

    def aux_param_shapes(self, input_shape):  # pylint: disable=unused-argument
        return {}


class Sequential(Container):
    """Feed-forward chain: the (tuple of) results of one module are the positional arguments of the next."""
    _module_name = 'sequential'

    def __init__(self, *modules, **kwargs):
        super(Sequential, self).__init__(kwargs.get('name'))
        for module in modules:
            if not isinstance(module, Module):
                raise TypeError('Sequential takes modules, got %r' % (module,))
        self._modules = list(modules)

    def _children(self):
        return self._modules

    def append(self, module):
        self._modules.append(module)

    def insert(self, index, module):
        self._modules.insert(index, module)

    def pop(self, index=-1):
        return self._modules.pop(index)

    def reverse(self):
        self._modules.reverse()

    def __iter__(self):
        return iter(self._modules)

    def __len__(self):
        return len(self._modules)

    def __getitem__(self, index):
        return self._modules[index]

    def __str__(self):
        return str(self._modules)

    def forward(self, *args):
        result = args
        for module in self._modules:
            result = module(*result)
            if not isinstance(result, tuple):
                result = (result,)
        return result[0] if len(result) == 1 else result


class Parallel(Container):
    """Modules applied to the same input."""


class Binary(Parallel):
    """`left(X) <op> right(X)`, what `module_a + module_b` builds (the residual connections of resnet.py)."""

    def __init__(self, left, right, op, name=None):
        super(Binary, self).__init__(name)
        self._left, self._right, self._operator = left, right, op

    def _children(self):
        return (self._left, self._right)

    def __str__(self):
        return '(%s %s %s)' % (self._left, self._module_name, self._right)

    def forward(self, X):
        return self._operator(self._left(X), self._right(X))


class Add(Binary):
    _module_name = 'add'

    def __init__(self, left, right, name=None):
        super(Add, self).__init__(left, right, operator.add, name)


class Sub(Binary):
    _module_name = 'sub'

    def __init__(self, left, right, name=None):
        super(Sub, self).__init__(left, right, operator.sub, name)


class Mul(Binary):
    _module_name = 'mul'

    def __init__(self, left, right, name=None):
        super(Mul, self).__init__(left, right, operator.mul, name)


def _register_configs(configs, to_register):
    """Merge `to_register` into `configs` ({parameter name: {attribute: value}}): a key that names a
    parameter carries a dict for that parameter only; any other key is an attribute set on EVERY
    parameter.  Parameter-specific values win over global ones."""
    specific = {}
    for key, value in to_register.items():
        if not isinstance(key, str):
            raise KeyError('config keys are parameter names or attribute names, got %r' % (key,))
        if key in configs:
            specific[key] = value
        else:
            for config in configs.values():
                config[key] = value
    for key, value in specific.items():
        configs[key].update(value)


class Layer(Module):
    """A module with parameters.  `params` / `aux_params` are the layer-local names; the model-wide
    names are `<layer name>_<local name>` and are available as attributes (`self.weight` is the
    model-wide NAME of the local parameter 'weight'; auxiliary ones as `self._moving_mean`)."""

    def __init__(self, params=None, aux_params=None, name=None):
        super(Layer, self).__init__(name)
        self._module_param_names = tuple(params or ())
        self._param_names = self._global_names(self._module_param_names)
        for local, full in zip(self._module_param_names, self._param_names):
            setattr(self, local, full)
        self._module_aux_param_names = tuple(aux_params or ())
        self._aux_param_names = self._global_names(self._module_aux_param_names)
        for local, full in zip(self._module_aux_param_names, self._aux_param_names):
            setattr(self, '_%s' % local, full)
        self._model = None
        self._init_configs = dict((full, {}) for full in self._param_names + self._aux_param_names)
        defaults = dict((local, self._get_default_init_config(local))
                        for local in self._module_param_names + self._module_aux_param_names)
        self._register_init_configs(defaults)
        self._update_configs = dict((full, {}) for full in self._param_names)
        self._register_update_configs({'update_rule': 'unspecified'})   # the Updater must name one
        self._mode = 'training'
        self._initialized = False

    def _global_names(self, local_names):
        return tuple('%s_%s' % (self._name, local) for local in local_names)

    # the reference's spelling
    _assign_param_names = lambda self, *names: self._global_names(names)

    def __call__(self, *args, **kwargs):
        if self._model is None:
            raise AssertionError('A layer must be bound to a model.')
        if not self._initialized:
            shapes = tuple(a.shape for a in args if _is_array(a))
            kwshapes = dict((k, v.shape) for k, v in kwargs.items() if _is_array(v))
            self._init_params(self.param_shapes(*shapes, **kwshapes))
            self._init_aux_params(self.aux_param_shapes(*shapes, **kwshapes))
            self._initialized = True
        return self.forward(*args, **kwargs)

    def _affiliate_to(self, model):
        model._update_configs.update(self._update_configs)  # pylint: disable=protected-access
        self._model = model

    @staticmethod
    def _get_default_init_config(param_name):
        if 'weight' in param_name:
            return {'init_rule': 'xavier'}
        if 'gamma' in param_name or 'moving_var' in param_name:
            return {'init_rule': 'constant', 'value': 1}
        return {'init_rule': 'constant', 'value': 0}      # bias, beta, moving_mean, anything else

    def _create(self, name, shape):
        config = self._init_configs[name]
        return getattr(minpy.nn.init, config['init_rule'])(tuple(shape), config)

    def _init_params(self, param_shapes):
        model = self._model
        for name, shape in param_shapes.items():
            if name not in model.params:            # pre-loaded parameters are kept
                model.params[name] = self._create(name, shape)
            if model._attach_all and model._is_recording:  # pylint: disable=protected-access
                model.attach(name, model.params[name])

    def _init_aux_params(self, aux_param_shapes):
        model = self._model
        for name, shape in aux_param_shapes.items():
            if name not in model.aux_params:
                model.aux_params[name] = self._create(name, shape)

    def _get_param(self, param_name):
        return self._model.params[param_name]

    def _get_params(self, *param_names):
        return tuple(self._model.params[name] for name in param_names)

    def _get_aux_param(self, aux_param_name):
        return self._model.aux_params[aux_param_name]

    def _get_aux_params(self, *aux_param_names):
        return tuple(self._model.aux_params[name] for name in aux_param_names)

    def _parse_param_configs(self, configs):
        """User configs -> {model-wide parameter name or attribute: value}: tuple keys fan out to each
        named parameter, layer-local parameter names become model-wide ones."""
        if configs is None:
            return {}
        out = {}
        for key, value in configs.items():
            for sub in (key if isinstance(key, tuple) else (key,)):
                out[sub] = value
        for local, full in zip(self._module_param_names + self._module_aux_param_names,
                               self._param_names + self._aux_param_names):
            if local in out:
                out[full] = out.pop(local)
        return out

    def _register_init_configs(self, init_configs):
        _register_configs(self._init_configs, self._parse_param_configs(init_configs))

    def _register_update_configs(self, update_configs):
        _register_configs(self._update_configs, self._parse_param_configs(update_configs))

    def training(self):
        self._mode = 'training'

    def inference(self):
        self._mode = 'inference'

    @property
    def param_dict(self):
        return dict(zip(self._param_names, self._get_params(*self._param_names)))

    @property
    def aux_param_dict(self):
        return dict(zip(self._aux_param_names, self._get_aux_params(*self._aux_param_names)))

    def param_shapes(self, *args, **kwargs):  # pylint: disable=unused-argument
        """{model-wide parameter name: shape} given the input shapes; layers with parameters override."""
        return {}

    def aux_param_shapes(self, *args, **kwargs):  # pylint: disable=unused-argument
        return {}


# -------------------------------------------------------------------------------- model ----
class Model(minpy.nn.model.ModelBase):
    """A `ModelBase` whose parameters are declared by the modules assigned to it."""

    def __init__(self, loss=None):
        object.__setattr__(self, '_modules', [])
        object.__setattr__(self, '_module_names', set())
        object.__setattr__(self, '_update_configs', {})
        super(Model, self).__init__()
        if loss is not None:
            self.loss = loss
        self._tape = None
        self._attach_all = True
        self._results = None
        self._bp_name_list, self._bp_array_list, self._bp_index_dict = [], [], {}

    @property
    def _is_recording(self):
        return self._tape is not None and self._tape.is_recording

    def __setattr__(self, attr, value):
        if isinstance(value, Model):
            self._register_model(value)
        elif isinstance(value, Module):
            self._register_module(value)
        elif isinstance(value, (tuple, list, set, frozenset)):
            self._register_iterable(value)
        object.__setattr__(self, attr, value)

    def _register_model(self, model):
        """A sub-model's parameters join this model's tape."""
        model.attach = lambda name, array: self.attach(name, array)
        model.detach = lambda name: self.detach(name)

    def _register_module(self, module):
        if any(module is known for known in self._modules):
            raise AssertionError('module %s is already registered' % module.name)
        if module.name in self._module_names:
            raise AssertionError('duplicate module name %s' % module.name)
        self._modules.append(module)
        self._module_names.add(module.name)
        module._affiliate_to(self)  # pylint: disable=protected-access

    def _register_iterable(self, iterable):
        for element in iterable:
            if isinstance(element, Module):
                self._register_module(element)
            elif isinstance(element, (tuple, list, set, frozenset)):
                self._register_iterable(element)

    # parameters come from layers, never from add_param
    def add_param(self, *args, **kwargs):
        raise NotImplementedError('parameters of a model_builder.Model are declared by its layers')

    add_params = add_aux_param = add_param

    def forward(self, *args, **kwargs):
        raise NotImplementedError()

    def forward_batch(self, batch, mode):
        return self.forward(batch.data[0], mode='training' if mode == 'train' else 'inference')

    def __call__(self, forward_args=None, forward_kwargs=None, labels=None, loss_kwargs=None, forward=None,
                 loss=None, attach_all=True, reduce_array=False):
        """Forward (and loss) under a fresh tape with the parameters marked; returns the loss (or the
        predictions when the model has no loss).  `loss` is a name in minpy.nn.layers or a callable
        `loss(*predictions, *labels, **loss_kwargs)`."""
        self._bp_name_list, self._bp_array_list, self._bp_index_dict = [], [], {}
        self._tape = minpy.tape.Tape()
        minpy.tape.Tape.global_tape = self._tape
        self._tape.start_recording()
        try:
            self._attach_all = attach_all
            if attach_all:
                for key, value in self.params.items():
                    self.attach(key, value)
            if forward_args is None:
                forward_args = ()
            elif not isinstance(forward_args, tuple):
                forward_args = (forward_args,)
            forward_args = tuple(minpy.array.wrap(a) for a in forward_args)
            forward_kwargs = dict(forward_kwargs or {})
            forward_kwargs['mode'] = 'training'
            predictions = (forward or self.forward)(*forward_args, **forward_kwargs)
            many = isinstance(predictions, (tuple, list))
            if not all(_is_minpy_value(p) for p in (predictions if many else (predictions,))):
                raise AssertionError('Forward function must return an array or an iterable of arrays.')
            if loss is None:
                loss = getattr(self, 'loss', None)
                if getattr(loss, '__func__', None) is minpy.nn.model.ModelBase.loss:
                    loss = None                       # the abstract method of ModelBase: no loss given
            if loss is None:
                self._results = predictions
            else:
                if isinstance(loss, str):
                    loss = getattr(minpy.nn.layers, loss)
                predictions = tuple(predictions) if many else (predictions,)
                labels = labels if isinstance(labels, tuple) else (labels,)
                labels = tuple(minpy.array.wrap(l) for l in labels)
                self._results = loss(*(predictions + labels), **(loss_kwargs or {}))
        finally:
            self._tape.stop_recording()
            minpy.tape.Tape.global_tape = None
        if not reduce_array:
            return self._results
        if isinstance(self._results, (tuple, list)):
            return [self._reduce_array(r) for r in self._results]
        return self._reduce_array(self._results)

    @staticmethod
    def _reduce_array(value):
        """A size-1 result as a Python float."""
        if _is_minpy_value(value) and functools.reduce(operator.mul, getattr(value, 'shape', ()), 1) == 1:
            return float(value)
        return value

    def backward(self, upstream=None):  # pylint: disable=unused-argument
        """{parameter name: gradient} of the last `__call__`'s result(s)."""
        if self._tape is None:
            raise RuntimeError('backward() needs a forward pass: call the model first')
        if isinstance(self._results, (tuple, list)):
            targets = tuple(minpy.array.wrap(r) for r in self._results)
        else:
            targets = minpy.array.wrap(self._results)
        grads = self._tape.get_gradient(tuple(self._bp_array_list), targets)
        return dict(zip(self._bp_name_list, grads))

    def attach(self, name, array):
        """Differentiate with respect to `array` under the name `name` (parameters are attached
        automatically unless the model was called with attach_all=False)."""
        if not self._is_recording:
            raise RuntimeError('attach() is only possible while the model is recording (inside __call__)')
        if name in self._bp_index_dict:
            return self
        array.mark_for_bp(self._tape)
        self._bp_index_dict[name] = len(self._bp_name_list)
        self._bp_name_list.append(name)
        self._bp_array_list.append(array)
        return self

    def detach(self, name):
        index = self._bp_index_dict.pop(name)
        del self._bp_name_list[index]
        self._bp_array_list[index]._bp_timestamp = -1  # pylint: disable=protected-access
        del self._bp_array_list[index]
        for key, value in self._bp_index_dict.items():
            if value > index:
                self._bp_index_dict[key] = value - 1
        return self

    def detach_graph(self):
        return self

    def grad_and_loss(self, data, labels):
        """(gradient dict, loss) of one batch."""
        loss = self(data, labels=labels)
        return self.backward(), loss

    def grad(self, data, labels):
        return self.grad_and_loss(data, labels)[0]

    def training(self):
        for module in self._modules:
            module.training()

    def inference(self):
        for module in self._modules:
            module.inference()


# ------------------------------------------------------------------------------ updater ----
class _ConfigParser(object):
    """Attribute view over {parameter name: config}: `parser.learning_rate = x` sets it for every
    parameter, `parser.learning_rate` reads it if all agree, `parser[name].learning_rate` is one
    parameter's."""

    class _ParamRef(object):
        def __init__(self, config):
            object.__setattr__(self, '_config', config)

        def __getattr__(self, attr):
            try:
                return self._config[attr]
            except KeyError:
                raise AttributeError(attr)

        def __setattr__(self, attr, value):
            self._config[attr] = value

    def __init__(self, configs):
        object.__setattr__(self, '_configs', configs)

    def __getattr__(self, attr):
        values = [config[attr] for config in self._configs.values() if attr in config]
        if not values or any(v is not values[0] and v != values[0] for v in values[1:]):
            raise AssertionError('Inconsistent or non-existent attribute.')
        return values[0]

    def __setattr__(self, attr, value):
        for config in self._configs.values():
            config[attr] = value

    def __getitem__(self, param_name):
        return _ConfigParser._ParamRef(self._configs[param_name])

    def __setitem__(self, param_name, config):
        assert isinstance(config, dict)
        self._configs[param_name] = copy.deepcopy(config)

    def keys(self):
        return self._configs.keys()

    def values(self):
        return self._configs.values()

    def items(self):
        return self._configs.items()


class Updater(_ConfigParser):
    """Applies nn/optim.py rules to a model's parameters from a gradient dict.  Keyword arguments of the
    constructor are set on every parameter (`update_rule='sgd'`, `learning_rate=0.1`, ...); several
    updaters of one model are independent (each copies the model's update configs)."""

    def __init__(self, model, **kwargs):
        super(Updater, self).__init__(copy.deepcopy(model._update_configs))  # pylint: disable=protected-access
        object.__setattr__(self, '_model', model)
        for attr, value in kwargs.items():
            setattr(self, attr, value)

    def _config_of(self, param_name):
        config = self._configs.get(param_name)
        if config is None:
            # a parameter created after this updater (layers initialise lazily): take the model's
            # config and this updater's global attributes
            config = copy.deepcopy(self._model._update_configs.get(param_name, {}))  # pylint: disable=protected-access
            for other in self._configs.values():
                for key, value in other.items():
                    if not _is_minpy_value(value) and not isinstance(value, numpy.ndarray) and key != 't':
                        config.setdefault(key, value)
                break
            self._configs[param_name] = config
        return config

    def __call__(self, grad_dict):
        """Update exactly the parameters named in `grad_dict`."""
        for param_name, grad in grad_dict.items():
            config = self._config_of(param_name)
            rule = config.get('update_rule', 'unspecified')
            if not hasattr(minpy.nn.optim, str(rule)):
                raise ValueError('update_rule of %s is %r: name one of nn/optim.py' % (param_name, rule))
            state = dict((k, v) for k, v in config.items() if k != 'update_rule')
            self._model.params[param_name], new_state = getattr(minpy.nn.optim, rule)(
                self._model.params[param_name], grad, state)
            config.update(new_state)
