"""The BASELINE.json training configurations as MinPy models, written the way
the reference's example scripts write theirs (examples/nn/mlp.py:23-60,
examples/nn/mlp_bn_dropout.py:26-60, examples/nn/lstm.py:9-60): a `ModelBase` subclass whose
`forward` / `loss` call `minpy.nn.layers`, trained by `minpy.nn.solver.Solver`.  Used by
bench.py and by the parity tests; also runnable against the NumPy oracle because each model
takes the layer namespace as an argument.
"""
from __future__ import division

import numpy as np

from minpy.nn import layers as _layers
from minpy.nn.model import ModelBase

CIFAR_SHAPE = (3, 32, 32)
CIFAR_DIM = 3 * 32 * 32


class MLP(ModelBase):
    """Config C1: affine(3072->512)-relu-affine(512->10), softmax loss (examples/nn/mlp.py)."""

    def __init__(self, in_dim=CIFAR_DIM, hidden=512, classes=10):
        super(MLP, self).__init__()
        self.in_dim = in_dim
        self.add_param(name='w1', shape=(in_dim, hidden)).add_param(name='b1', shape=(hidden,)) \
            .add_param(name='w2', shape=(hidden, classes)).add_param(name='b2', shape=(classes,))

    def forward(self, X, mode):
        X = X.reshape((X.shape[0], self.in_dim))
        h = _layers.relu(_layers.affine(X, self.params['w1'], self.params['b1']))
        return _layers.affine(h, self.params['w2'], self.params['b2'])

    def loss(self, predict, y):
        return _layers.softmax_loss(predict, y)


class DeepMLP(ModelBase):
    """Config C3: 3072-512-512-512-512-10 with batchnorm + dropout(0.5) after every hidden
    relu (the block of examples/nn/mlp_bn_dropout.py:41-54, repeated)."""

    def __init__(self, in_dim=CIFAR_DIM, hidden=(512, 512, 512, 512), classes=10, drop=0.5):
        super(DeepMLP, self).__init__()
        self.in_dim, self.drop = in_dim, drop
        dims = [in_dim] + list(hidden)
        self.depth = len(hidden)
        for i in range(self.depth):
            self.add_param(name='w%d' % i, shape=(dims[i], dims[i + 1])) \
                .add_param(name='b%d' % i, shape=(dims[i + 1],)) \
                .add_param(name='gamma%d' % i, shape=(dims[i + 1],), init_rule='constant', init_config={'value': 1.0}) \
                .add_param(name='beta%d' % i, shape=(dims[i + 1],), init_rule='constant') \
                .add_aux_param(name='running_mean%d' % i, value=None) \
                .add_aux_param(name='running_var%d' % i, value=None)
        self.add_param(name='wo', shape=(dims[-1], classes)).add_param(name='bo', shape=(classes,))

    def forward(self, X, mode):
        h = X.reshape((X.shape[0], self.in_dim))
        for i in range(self.depth):
            h = _layers.relu(_layers.affine(h, self.params['w%d' % i], self.params['b%d' % i]))
            h, self.aux_params['running_mean%d' % i], self.aux_params['running_var%d' % i] = _layers.batchnorm(
                h, self.params['gamma%d' % i], self.params['beta%d' % i], mode=mode,
                running_mean=self.aux_params['running_mean%d' % i], running_var=self.aux_params['running_var%d' % i])
            h = _layers.dropout(h, self.drop, mode=mode)
        return _layers.affine(h, self.params['wo'], self.params['bo'])

    def loss(self, predict, y):
        return _layers.softmax_loss(predict, y)


class CharLSTM(ModelBase):
    """Config C5: one-hot tokens -> LSTM(H) over T steps -> per-step vocabulary logits,
    masked temporal softmax loss (examples/nn/lstm.py:9-60 with a temporal read-out)."""

    def __init__(self, vocab=128, hidden=1024):
        super(CharLSTM, self).__init__()
        self.V, self.H = vocab, hidden
        self.add_param(name='Wx', shape=(vocab, 4 * hidden)).add_param(name='Wh', shape=(hidden, 4 * hidden)) \
            .add_param(name='b', shape=(4 * hidden,), init_rule='constant') \
            .add_param(name='Wa', shape=(hidden, vocab)).add_param(name='ba', shape=(vocab,), init_rule='constant')

    def forward(self, X, mode):
        import minpy.numpy as mnp
        N = X.shape[0]
        h0 = mnp.zeros((N, self.H))
        hs = _layers.lstm_temporal(X, h0, self.params['Wx'], self.params['Wh'], self.params['b'])
        return _layers.temporal_affine(hs, self.params['Wa'], self.params['ba'])

    def loss(self, predict, y):
        import minpy.numpy as mnp
        mask = mnp.ones((y.shape[0], y.shape[1]))
        return _layers.temporal_softmax_loss(predict, y, mask)


class ConvNet(ModelBase):
    """Config C4: 3 x (conv3x3 pad 1 -> relu -> maxpool 2/2) with 32/64/128 filters, then
    affine 2048 -> 10.  Written as examples/nn/cnn.py:26-60 writes its network: the
    convolutional trunk is an `mx.sym` graph wrapped in `minpy.core.Function`, the classifier
    is `nn.layers`."""

    def __init__(self, batch=512, filters=(32, 64, 128), classes=10):
        super(ConvNet, self).__init__()
        import mxnet as mx
        from minpy.core import Function
        net = mx.sym.Variable(name='X')
        for i, f in enumerate(filters):
            net = mx.sym.Convolution(data=net, name='conv%d' % i, kernel=(3, 3), pad=(1, 1), num_filter=f)
            net = mx.sym.Activation(data=net, act_type='relu')
            net = mx.sym.Pooling(data=net, name='pool%d' % i, pool_type='max', kernel=(2, 2), stride=(2, 2))
        net = mx.sym.Flatten(data=net)
        self.conv = Function(net, input_shapes={'X': (batch,) + CIFAR_SHAPE}, name='conv')
        self.add_params(self.conv.get_params())
        feat = self.conv.get_one_output_shape()[1]
        self.add_param(name='w1', shape=(feat, classes)).add_param(name='b1', shape=(classes,))

    def forward(self, X, mode):
        self.conv.is_train = (mode == 'train')
        out = self.conv(X=X, **self.params)
        return _layers.affine(out, self.params['w1'], self.params['b1'])

    def loss(self, predict, y):
        return _layers.softmax_loss(predict, y)


class RefConvNet(ModelBase):
    """The network of examples/nn/cnn.py:26-60 itself: conv7x7x32 -> relu -> maxpool 2/2 ->
    affine 5408 -> 512 -> relu -> affine 512 -> 10, batch 128."""

    def __init__(self, batch=128, hidden=512, classes=10):
        super(RefConvNet, self).__init__()
        import mxnet as mx
        from minpy.core import Function
        net = mx.sym.Variable(name='X')
        net = mx.sym.Convolution(data=net, name='conv', kernel=(7, 7), num_filter=32)
        net = mx.sym.Activation(data=net, act_type='relu')
        net = mx.sym.Pooling(data=net, name='pool', pool_type='max', kernel=(2, 2), stride=(2, 2))
        net = mx.sym.Flatten(data=net)
        self.conv = Function(net, input_shapes={'X': (batch,) + CIFAR_SHAPE}, name='conv')
        self.add_params(self.conv.get_params())
        feat = self.conv.get_one_output_shape()[1]
        self.add_param(name='w1', shape=(feat, hidden)).add_param(name='b1', shape=(hidden,)) \
            .add_param(name='w2', shape=(hidden, classes)).add_param(name='b2', shape=(classes,))

    def forward(self, X, mode):
        self.conv.is_train = (mode == 'train')
        out = self.conv(X=X, **self.params)
        out = _layers.relu(_layers.affine(out, self.params['w1'], self.params['b1']))
        return _layers.affine(out, self.params['w2'], self.params['b2'])

    def loss(self, predict, y):
        return _layers.softmax_loss(predict, y)


# ------------------------------------------------------------------------ synthetic data --
def cifar_like(num, seed=0, dtype=np.float32):
    """X ~ N(0,1) of shape (num, 3, 32, 32), y ~ U{0..9}."""
    x = np.random.RandomState(seed).randn(num, *CIFAR_SHAPE).astype(dtype)
    y = np.random.RandomState(seed + 1).randint(0, 10, num).astype(np.int32)
    return x, y


def char_like(num, T=100, V=128, seed=5, dtype=np.float32):
    tok = np.random.RandomState(seed).randint(0, V, (num, T))
    x = np.eye(V, dtype=dtype)[tok]
    y = np.random.RandomState(seed + 1).randint(0, V, (num, T)).astype(np.int32)
    return x, y


def init_params(model, seed, scale=0.01):
    """name -> float32 host array: weights scale*N(0,1), biases/beta 0, gamma 1."""
    rng = np.random.RandomState(seed)
    out = {}
    for name, cfg in model.param_configs.items():
        shape = cfg['shape']
        if name.startswith('gamma'):
            out[name] = np.ones(shape, np.float32)
        elif len(shape) == 1:
            out[name] = np.zeros(shape, np.float32)
        else:
            out[name] = (rng.randn(*shape) * scale).astype(np.float32)
    return out


WORKLOADS = {
    # name: (model factory, per-GPU batch, update rule, optim config, data maker, description)
    'mlp_c1': (MLP, 256, 'sgd_momentum', {'learning_rate': 1e-4, 'momentum': 0.9}, cifar_like,
               '2-layer MLP 3072-512-10, batch 256, sgd_momentum'),
    'mlp_c3': (DeepMLP, 1024, 'sgd_momentum', {'learning_rate': 1e-3, 'momentum': 0.9}, cifar_like,
               '5-layer MLP 3072-512x4-10 with batchnorm+dropout, batch 1024, sgd_momentum'),
    'cnn_c4': (ConvNet, 512, 'sgd_momentum', {'learning_rate': 1e-3, 'momentum': 0.9}, cifar_like,
               'CNN 3x(conv3x3-relu-maxpool) 32/64/128 + affine 2048-10, batch 512, sgd_momentum'),
    'cnn_ref': (RefConvNet, 128, 'sgd_momentum', {'learning_rate': 1e-3, 'momentum': 0.9}, cifar_like,
                'examples/nn/cnn.py net: conv7x7x32-relu-maxpool + affine 5408-512-10, batch 128, sgd_momentum'),
    'lstm_c5': (CharLSTM, 128, 'adam', {'learning_rate': 1e-3}, char_like,
                'char LSTM V=128 H=1024 T=100, batch 128, adam'),
}
