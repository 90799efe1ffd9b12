"""First-order update rules: `rule(w, dw, config) -> (next_w, config)`.

Formulas and config keys are those of the reference's minpy/nn/optim.py:33-141 (itself after
cs231n); state lives in the per-parameter `config` dict.  Under a CUDA policy each rule is one
fused kernel per parameter tensor instead of 4-12 elementwise passes.
"""
import minpy.numpy as np
from minpy.array_variants import ArrayType

# pylint: disable=no-member


def _cuda(name):
    prim = getattr(np, name, None)
    return prim if getattr(prim, 'type', None) == ArrayType.CUDA else None


def sgd(w, dw, config=None):
    """w - learning_rate * dw."""
    if config is None:
        config = {}
    config.setdefault('learning_rate', 1e-2)
    fused = _cuda('fused_sgd')
    if fused is not None:
        return fused(w, dw, config['learning_rate']), config
    w -= config['learning_rate'] * dw
    return w, config


def sgd_momentum(w, dw, config=None):
    """v = momentum * v - learning_rate * dw; w + v.  State: `velocity`."""
    if config is None:
        config = {}
    config.setdefault('learning_rate', 1e-2)
    config.setdefault('momentum', 0.9)
    v = config.get('velocity')
    if v is None:
        v = np.zeros_like(w)
    fused = _cuda('fused_sgd_momentum')
    if fused is not None:
        next_w, v = fused(w, dw, v, config['learning_rate'], config['momentum'])
    else:
        v = v * config['momentum'] - dw * config['learning_rate']
        next_w = w + v
    config['velocity'] = v
    return next_w, config


def rmsprop(x, dx, config=None):
    """cache = decay * cache + (1 - decay) * dx^2; x - lr * dx / (sqrt(cache) + eps)."""
    if config is None:
        config = {}
    config.setdefault('learning_rate', 1e-2)
    config.setdefault('decay_rate', 0.99)
    config.setdefault('epsilon', 1e-8)
    if 'cache' not in config:
        config['cache'] = np.zeros_like(x)
    cache = config['cache']
    fused = _cuda('fused_rmsprop')
    if fused is not None:
        next_x, cache = fused(x, dx, cache, config['learning_rate'], config['decay_rate'], config['epsilon'])
    else:
        cache = cache * config['decay_rate'] + dx ** 2 * (1 - config['decay_rate'])
        next_x = x - config['learning_rate'] * dx / (np.sqrt(cache) + config['epsilon'])
    config['cache'] = cache
    return next_x, config


def adam(x, dx, config=None):
    """Adam with bias-corrected moments; epsilon is added outside the square root."""
    if config is None:
        config = {}
    config.setdefault('learning_rate', 1e-3)
    config.setdefault('beta1', 0.9)
    config.setdefault('beta2', 0.999)
    config.setdefault('epsilon', 1e-8)
    if 'm' not in config:
        config['m'] = np.zeros_like(x)
    if 'v' not in config:
        config['v'] = np.zeros_like(x)
    config.setdefault('t', 0)
    m, v, t = config['m'], config['v'], config['t'] + 1
    fused = _cuda('fused_adam')
    if fused is not None:
        next_x, m, v = fused(x, dx, m, v, config['learning_rate'], config['beta1'], config['beta2'],
                             config['epsilon'], t)
    else:
        m = config['beta1'] * m + (1 - config['beta1']) * dx
        v = config['beta2'] * v + (1 - config['beta2']) * (dx ** 2)
        m_hat = m / (1 - config['beta1'] ** t)
        v_hat = v / (1 - config['beta2'] ** t)
        next_x = x - config['learning_rate'] * m_hat / (np.sqrt(v_hat) + config['epsilon'])
    config['m'], config['v'], config['t'] = m, v, t
    return next_x, config
