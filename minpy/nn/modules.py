"""Ready-made layers for `minpy.nn.model_builder` (reference: minpy/nn/modules.py:12-401).

Two families, as in the reference:
  * plain layers written with `minpy.numpy` / `minpy.nn.layers` (Variable, Identity, ReLU, Dropout,
    Logistic, Tanh, Reshape, BatchReshape, Flatten, BatchFlatten, Embedding, RNN, LSTM);
  * `Symbolic` layers that wrap a one-operator `mx.sym` graph in `minpy.core.Function`
    (FullyConnected, Convolution, Pooling, BatchNorm) -- here the graph is the façade of
    minpy_b200/symbol.py and runs on the sm_100a kernels through minpy_b200/symbol_exec.py.
    Parameter names follow MXNet's (`<layer>_weight`, `_bias`, `_gamma`, `_beta`, auxiliary
    `_moving_mean` / `_moving_var`), FullyConnected weights are (num_hidden, in).

Departures from the reference: `Dropout.forward` works (the reference's body refers to undefined names,
modules.py:60-61) and honours training / inference mode; RNN / LSTM are written on `layers.rnn_step` /
`layers.lstm_step` (parameters Wx, Wh, b) instead of a SliceChannel / Group symbol graph, which the façade
does not restate -- same call signature and results, different parameter names.
"""
from __future__ import absolute_import

import functools
import operator

import minpy.core
import minpy.numpy as np
import minpy.nn.layers
from minpy.nn.model_builder import Layer

# pylint: disable=no-member, arguments-differ


def _prod(shape):
    return functools.reduce(operator.mul, shape, 1)


class Variable(Layer):
    """A free parameter of the given shape; calling the layer returns it."""
    _module_name = 'variable'

    def __init__(self, shape, init_configs=None, update_configs=None, name=None):
        self._shape = tuple(shape)
        super(Variable, self).__init__(('variable',), None, name)
        self._register_init_configs(init_configs)
        self._register_update_configs(update_configs)

    def forward(self):
        return self._get_param(self.variable)

    def param_shapes(self, *args):
        return {self.variable: self._shape}


class Identity(Layer):
    _module_name = 'identity'

    def __init__(self, name=None):
        super(Identity, self).__init__(name=name)

    def forward(self, X):
        return X


class ReLU(Layer):
    _module_name = 'ReLU'

    def __init__(self, name=None):
        super(ReLU, self).__init__(name=name)

    def forward(self, X, *args):
        return minpy.nn.layers.relu(X)


class Dropout(Layer):
    """Zero each activation with probability p while training (survivors are not rescaled; inference
    scales by 1 - p: nn/layers.py:121-147)."""
    _module_name = 'dropout'

    def __init__(self, p, name=None):
        super(Dropout, self).__init__(name=name)
        self._p = p

    def forward(self, data):
        return minpy.nn.layers.dropout(data, self._p, mode='train' if self._mode == 'training' else 'test')


class Logistic(Layer):
    _module_name = 'logistic'

    def __init__(self, name=None):
        super(Logistic, self).__init__(name=name)

    def forward(self, X, *args):
        return minpy.nn.layers.sigmoid(X)


class Tanh(Layer):
    _module_name = 'tanh'

    def __init__(self, name=None):
        super(Tanh, self).__init__(name=name)

    def forward(self, X, *args):
        return np.tanh(X)


class Reshape(Layer):
    _module_name = 'reshape'

    def __init__(self, shape, name=None):
        super(Reshape, self).__init__(name=name)
        self._shape = tuple(shape)

    def forward(self, X, *args):
        return np.reshape(X, self._shape)


class BatchReshape(Layer):
    """Reshape everything but the batch axis."""
    _module_name = 'batch_reshape'

    def __init__(self, shape, name=None):
        super(BatchReshape, self).__init__(name=name)
        self._shape = tuple(shape)

    def forward(self, X, *args):
        return np.reshape(X, X.shape[:1] + self._shape)


class Flatten(Layer):
    _module_name = 'flatten'

    def __init__(self, name=None):
        super(Flatten, self).__init__(name=name)

    def forward(self, X):
        return np.reshape(X, (_prod(X.shape),))


class BatchFlatten(Layer):
    _module_name = 'batch_flatten'

    def __init__(self, name=None):
        super(BatchFlatten, self).__init__(name=name)

    def forward(self, X):
        return np.reshape(X, (X.shape[0], _prod(X.shape[1:])))


class Embedding(Layer):
    """Rows of a (input_dim, output_dim) table selected by integer indices."""
    _module_name = 'embedding'

    def __init__(self, input_dim, output_dim, init_configs=None, update_configs=None, name=None):
        super(Embedding, self).__init__(('weight',), None, name)
        self._dims = (int(input_dim), int(output_dim))
        self._register_init_configs(init_configs)
        self._register_update_configs(update_configs)

    def forward(self, data):
        return self._get_param(self.weight)[data]

    def param_shapes(self, input_shape):
        return {self.weight: self._dims}


# ------------------------------------------------------------------------- symbolic layers ----
class Symbolic(Layer):
    """A symbol graph as a layer: its Variables other than `variables` (default: 'data') are the layer's
    parameters, its auxiliary states the layer's auxiliary parameters."""
    _module_name = 'symbolic'

    def __init__(self, symbol, variables=None, init_configs=None, update_configs=None, name=None):
        self._symbol = symbol
        self._variables = ('data',) if variables is None else tuple(variables)
        params = tuple(a for a in symbol.list_arguments() if a not in self._variables)
        super(Symbolic, self).__init__(params, tuple(symbol.list_auxiliary_states()), name)
        self._register_init_configs(init_configs)
        self._register_update_configs(update_configs)
        self._func = None

    def forward(self, **kwargs):
        """kwargs: every variable of the graph by name."""
        if self._func is None:
            shapes = dict((key, value.shape) for key, value in kwargs.items())
            self._func = minpy.core.Function(self._symbol, shapes, name=self._name)
        kwargs.update(zip(self._module_param_names, self._get_params(*self._param_names)))
        kwargs.update(zip(self._module_aux_param_names, self._get_aux_params(*self._aux_param_names)))
        self._func.is_train = self._mode == 'training'
        return self._func(**kwargs)

    def param_shapes(self, **kwargs):
        shapes, _, _ = self._symbol.infer_shape(**kwargs)
        by_local = dict(zip(self._symbol.list_arguments(), shapes))
        return dict((full, tuple(by_local[local])) for local, full in zip(self._module_param_names, self._param_names))

    def aux_param_shapes(self, **kwargs):
        _, _, shapes = self._symbol.infer_shape(**kwargs)
        return dict(zip(self._aux_param_names, (tuple(s) for s in shapes)))


def _one_op(op_name, args, kwargs):
    """`mx.sym.<op>(data, *args, **kwargs)` without the keys that configure the layer, not the operator."""
    import mxnet
    kw = dict((k, v) for k, v in kwargs.items() if k not in ('init_configs', 'update_configs', 'name'))
    return getattr(mxnet.symbol, op_name)(mxnet.symbol.Variable('data'), *args, name=op_name.lower(), **kw)


class _DataSymbolic(Symbolic):
    """Symbolic layer of one operator with the single input 'data'."""
    _op = None
    _has_aux = False

    def __init__(self, *args, **kwargs):
        super(_DataSymbolic, self).__init__(_one_op(self._op, args, kwargs), None, kwargs.get('init_configs'),
                                            kwargs.get('update_configs'), kwargs.get('name'))

    def forward(self, data):
        return super(_DataSymbolic, self).forward(data=data)

    def param_shapes(self, input_shape):
        return super(_DataSymbolic, self).param_shapes(data=input_shape)

    def aux_param_shapes(self, input_shape):
        return super(_DataSymbolic, self).aux_param_shapes(data=input_shape) if self._has_aux else {}


class FullyConnected(_DataSymbolic):
    """`FullyConnected(num_hidden=..., no_bias=False)`: data (N, ...) . weight(num_hidden, in)^T + bias."""
    _module_name = 'fully_connected'
    _op = 'FullyConnected'


class Convolution(_DataSymbolic):
    """`Convolution(num_filter=, kernel=, stride=(1, 1), pad=(0, 0), no_bias=False)` on NCHW data."""
    _module_name = 'convolution'
    _op = 'Convolution'


class Pooling(_DataSymbolic):
    """`Pooling(pool_type='max'|'avg', kernel=, stride=, global_pool=False)` on NCHW data."""
    _module_name = 'pooling'
    _op = 'Pooling'


class BatchNorm(_DataSymbolic):
    """`BatchNorm(eps=1e-3, momentum=0.9, fix_gamma=True)` over axis 1 of (N, C) or (N, C, H, W) data;
    the moving statistics are auxiliary parameters of the model."""
    _module_name = 'batch_norm'
    _op = 'BatchNorm'
    _has_aux = True


# --------------------------------------------------------------------------- recurrent cells ----
class RNN(Layer):
    """One step of an Elman RNN: act(data . Wx + hidden . Wh + b); `hidden=None` starts from zeros."""
    _module_name = 'rnn'

    def __init__(self, num_hidden, act_type='tanh', init_configs=None, update_configs=None, name=None):
        if act_type not in ('tanh', 'relu', 'sigmoid'):
            raise ValueError('act_type must be tanh, relu or sigmoid')
        super(RNN, self).__init__(('Wx', 'Wh', 'b'), None, name)
        self._num_hidden, self._act_type = int(num_hidden), act_type
        self._register_init_configs({'Wx': {'init_rule': 'xavier'}, 'Wh': {'init_rule': 'xavier'}})
        self._register_init_configs(init_configs)
        self._register_update_configs(update_configs)

    def forward(self, data, hidden=None):
        if hidden is None:
            hidden = np.zeros((data.shape[0], self._num_hidden))
        Wx, Wh, b = self._get_params(self.Wx, self.Wh, self.b)
        if self._act_type == 'tanh':
            return minpy.nn.layers.rnn_step(data, hidden, Wx, Wh, b)
        pre = np.dot(data, Wx) + np.dot(hidden, Wh) + b
        return minpy.nn.layers.relu(pre) if self._act_type == 'relu' else minpy.nn.layers.sigmoid(pre)

    def param_shapes(self, data_shape, hidden_shape=None):
        H = self._num_hidden
        return {self.Wx: (data_shape[1], H), self.Wh: (H, H), self.b: (H,)}


class LSTM(Layer):
    """One LSTM step (nn/layers.py:353-380): returns (next_hidden, next_cell)."""
    _module_name = 'lstm'

    def __init__(self, num_hidden, act_type='tanh', init_configs=None, update_configs=None, name=None):
        if act_type != 'tanh':
            raise ValueError('the lstm_step kernel implements the tanh cell')
        super(LSTM, self).__init__(('Wx', 'Wh', 'b'), None, name)
        self._num_hidden = int(num_hidden)
        self._register_init_configs({'Wx': {'init_rule': 'xavier'}, 'Wh': {'init_rule': 'xavier'}})
        self._register_init_configs(init_configs)
        self._register_update_configs(update_configs)

    def forward(self, data, hidden=None, cell=None):
        N, H = data.shape[0], self._num_hidden
        if hidden is None:
            hidden = np.zeros((N, H))
        if cell is None:
            cell = np.zeros((N, H))
        return minpy.nn.layers.lstm_step(data, hidden, cell, *self._get_params(self.Wx, self.Wh, self.b))

    def param_shapes(self, data_shape, hidden_shape=None, cell_shape=None):
        H = self._num_hidden
        return {self.Wx: (data_shape[1], 4 * H), self.Wh: (H, 4 * H), self.b: (4 * H,)}
