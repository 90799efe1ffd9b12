"""Functional neural-network layers.

Every function has the signature and arithmetic of the reference's minpy/nn/layers.py:10-480.
Each one exists in two forms: the primitive composition (what the reference wrote; it is what
runs under the NumPy policy and therefore what the oracle computes), and, when the bound
namespace offers a CUDA `fused_*` primitive, a single fused kernel pair that computes the same
forward value and the same gradients (minpy/array_variants/cuda/cuda_core.py).
"""
from __future__ import division

import minpy.numpy as np
import minpy.numpy.random as random
from minpy.array_variants import ArrayType

# pylint: disable=invalid-name, too-many-arguments, too-many-locals, no-member


def _cuda(name):
    """The primitive bound to `name` if it is a CUDA one, else None."""
    prim = getattr(np, name, None)
    return prim if getattr(prim, 'type', None) == ArrayType.CUDA else None


def affine(x, w, b):
    """out = x.w + b for x (N, D), w (D, M), b (M,)."""
    fused = _cuda('fused_affine')
    if fused is not None and len(x.shape) == 2 and len(w.shape) == 2:
        return fused(x, w, b)
    return np.dot(x, w) + b


def relu(x):
    """max(0, x); the gradient is passed where x >= 0."""
    fused = _cuda('fused_relu')
    if fused is not None:
        return fused(x)
    return np.maximum(0, x)


def batchnorm(x, gamma, beta, mode='train', eps=1e-5, momentum=0.9, running_mean=None, running_var=None):
    """Batch normalisation over axis 0 of x (N, D).

    Train mode normalises with the batch mean and the biased batch variance and returns the
    exponentially averaged statistics `momentum * running + (1 - momentum) * batch`; test mode
    normalises with the running statistics.  Returns (out, running_mean, running_var).
    """
    N, D = x.shape
    N, D = int(N), int(D)
    if running_mean is None:
        running_mean = np.zeros(D)
    if running_var is None:
        running_var = np.zeros(D)
    if mode == 'train':
        fused = _cuda('fused_batchnorm')
        if fused is not None:
            return fused(x, gamma, beta, running_mean, running_var, eps, momentum)
        mean = np.sum(x, axis=0) / N
        centred = x - np.expand_dims(mean, axis=0)
        var = np.sum(centred ** 2, axis=0) / N
        inv_std = 1.0 / np.sqrt(var + eps)
        x_hat = centred * np.expand_dims(inv_std, axis=0)
        out = gamma * x_hat + beta
        running_mean = momentum * running_mean + (1.0 - momentum) * mean
        running_var = momentum * running_var + (1.0 - momentum) * var
    elif mode == 'test':
        fused = _cuda('fused_batchnorm_test')
        if fused is not None:
            return fused(x, gamma, beta, running_mean, running_var, eps), running_mean, running_var
        x_hat = (x - running_mean) / np.sqrt(running_var + eps)
        out = gamma * x_hat + beta
    else:
        raise ValueError('Invalid forward batchnorm mode "%s"' % mode)
    return out, running_mean, running_var


def dropout(x, prob, mode='train', seed=None):
    """Train: zero each element with probability `prob` (survivors are NOT rescaled); any other
    mode: scale by (1 - prob).  `seed` makes the mask reproducible."""
    if seed is not None:
        random.seed(seed)
    if mode == 'train':
        fused = _cuda('fused_dropout')
        if fused is not None:
            from minpy.array_variants.cuda.random import next_stream
            stream_seed, offset = next_stream(x.size)
            return fused(x, prob, stream_seed, offset)
        mask = random.rand(*x.shape) > prob
        return x * mask
    return x * (1 - prob)


def svm_loss(x, y):
    """Multiclass hinge loss (margin 1) of scores x (N, C) against labels y (N,)."""
    N = x.shape[0]
    rows = np.arange(N)
    correct = x[rows, y]
    margins = np.maximum(0, x - correct + 1.0)
    return (np.sum(margins) - np.sum(margins[rows, y])) / N


def _as_onehot(label, N, C):
    if len(label.shape) == 1:
        onehot = np.zeros([N, C])
        np.onehot_encode(label, onehot)
        return onehot
    return label


def softmax_cross_entropy(prob, label):
    """Mean cross entropy of probabilities (N, C) against index (N,) or one-hot (N, C) labels."""
    N, C = prob.shape[0], prob.shape[1]
    onehot = _as_onehot(label, N, C)
    return -np.sum(np.log(prob) * onehot) / N


def softmax_loss(x, label):
    """Mean softmax cross entropy of scores x (N, C); gradient w.r.t. x is (softmax - y) / N."""
    fused = _cuda('fused_softmax_loss')
    if fused is not None:
        return fused(x, label)
    N, C = x.shape[0], x.shape[1]
    onehot = _as_onehot(label, N, C)
    prob = np.softmax_output(x, onehot)
    return softmax_cross_entropy(prob, onehot)


def l2_loss(x, label):
    """Mean (over the batch) squared error against index or one-hot labels."""
    N, C = x.shape[0], x.shape[1]
    onehot = _as_onehot(label, N, C)
    return np.sum((x - onehot) ** 2) / N


def sigmoid(x):
    """Logistic function."""
    fused = _cuda('sigmoid')
    if fused is not None:
        return fused(x)
    return 1 / (1 + np.exp(-x))


def rnn_step(x, prev_h, Wx, Wh, b):
    """Vanilla tanh RNN cell: x (N, D), prev_h (N, H), Wx (D, H), Wh (H, H), b (H,)."""
    return np.tanh(np.dot(x, Wx) + np.dot(prev_h, Wh) + b)


def rnn_temporal(x, h0, Wx, Wh, b):
    """Run `rnn_step` over x (N, T, D); returns the hidden states (N, T, H)."""
    N, T, _ = x.shape
    H = h0.shape[1]
    h = np.zeros([N, 0, H])
    prev = h0
    for t in range(T):
        prev = rnn_step(x[:, t, :], prev, Wx, Wh, b)
        h = np.append(h, prev.reshape(N, 1, H), axis=1)
    return h


def gru_step(x, prev_h, Wx, Wh, b, Wxh, Whh, bh):
    """GRU cell (Jozefowicz et al. 2015 variant): Wx (D, 2H), Wh (H, 2H), b (2H,) drive the
    reset/update gates; Wxh (D, H), Whh (H, H), bh (H,) the candidate state."""
    _, H = prev_h.shape
    gates = sigmoid(np.dot(x, Wx) + np.dot(prev_h, Wh) + b)
    r = gates[:, 0:H]
    z = gates[:, H:2 * H]
    candidate = np.tanh(np.dot(x, Wxh) + np.dot(r * prev_h, Whh) + bh)
    return z * prev_h + (1 - z) * candidate


def lstm_step(x, prev_h, prev_c, Wx, Wh, b):
    """LSTM cell with gate order (input, forget, output, candidate) along the 4H axis.
    Returns (next_h, next_c)."""
    fused = _cuda('fused_lstm_step')
    if fused is not None:
        return fused(x, prev_h, prev_c, Wx, Wh, b)
    _, H = prev_c.shape
    a = np.dot(x, Wx) + np.dot(prev_h, Wh) + b
    i = sigmoid(a[:, 0:H])
    f = sigmoid(a[:, H:2 * H])
    o = sigmoid(a[:, 2 * H:3 * H])
    g = np.tanh(a[:, 3 * H:4 * H])
    next_c = f * prev_c + i * g
    next_h = o * np.tanh(next_c)
    return next_h, next_c


def lstm_temporal(x, h0, Wx, Wh, b):
    """Run `lstm_step` over x (N, T, D) from hidden state h0 and a zero cell state; returns the
    hidden states (N, T, H)."""
    fused = _cuda('fused_lstm_temporal')
    if fused is not None:
        return fused(x, h0, Wx, Wh, b)
    N, T, _ = x.shape
    _, H = h0.shape
    h = np.zeros([N, 0, H])
    prev_h, prev_c = h0, np.zeros((N, H))
    for t in range(T):
        prev_h, prev_c = lstm_step(x[:, t, :], prev_h, prev_c, Wx, Wh, b)
        h = np.append(h, prev_h.reshape(N, 1, H), axis=1)
    return h


def temporal_affine(x, w, b):
    """Affine map applied to every timestep: x (N, T, D), w (D, M), b (M,) -> (N, T, M)."""
    N, T, D = x.shape
    M = b.shape[0]
    fused = _cuda('fused_affine')
    if fused is not None:        # bias in the GEMM epilogue instead of a broadcast add (and its column-sum VJP pass)
        return fused(x.reshape(N * T, D), w, b).reshape(N, T, M)
    return x.reshape(N * T, D).dot(w).reshape(N, T, M) + b


def temporal_softmax_loss(x, y, mask):
    """Masked softmax cross entropy over scores x (N, T, V), targets y (N, T) and mask (N, T):
    summed over time, averaged over the batch."""
    fused = _cuda('fused_temporal_softmax_loss')
    if fused is not None and len(x.shape) == 3:
        return fused(x, y, mask)
    N, T, V = x.shape
    x_flat = x.reshape(N * T, V)
    y_flat = y.reshape(N * T)
    mask_flat = mask.reshape(N * T)
    probs = np.exp(x_flat - np.max(x_flat, axis=1, keepdims=True))
    probs = probs / np.sum(probs, axis=1, keepdims=True)
    return -np.sum(mask_flat * np.log(probs[np.arange(N * T), y_flat])) / N
