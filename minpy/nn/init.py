"""Parameter initialisers `rule(shape, config) -> Array` (reference: minpy/nn/init.py:9-103)."""
import numpy

import minpy.numpy as np
import minpy.numpy.random as npr

# pylint: disable=no-member


def xavier(shape, _):
    """Normal draws scaled by sqrt(6 / (shape[0] + prod(shape[1:]))) -- normal, not uniform,
    and with shape[0] counted as fan-out: both as in the reference (init.py:27-34)."""
    fan_out = shape[0]
    fan_in = numpy.prod(shape[1:]) if len(shape) > 1 else 0
    scale = numpy.sqrt(6.0 / (fan_out + fan_in))
    return npr.randn(*shape) * scale


def constant(shape, config):
    """Every element equal to config['value'] (default 0)."""
    config.setdefault('value', 0.0)
    return np.ones(shape) * config['value']


def gaussian(shape, config):
    """Normal with mean config['mu'] (default 0) and std config['stdvar'] (default 1e-3)."""
    config.setdefault('mu', 0.0)
    config.setdefault('stdvar', 0.001)
    return npr.randn(*shape) * config['stdvar'] + config['mu']


def custom(shape, config):
    """config['function'](shape); zeros when no function is given."""
    return config.setdefault('function', np.zeros)(shape)
