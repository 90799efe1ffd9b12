"""Random-direction central-difference gradient check (reference:
minpy/utils/gradient_checker.py:11-62)."""
from __future__ import absolute_import
from __future__ import print_function

import numpy as np

from minpy.array import Value
from minpy.core import grad


def quick_grad_check(fun, arg, verbose=True, eps=1e-2, rtol=1e-2, atol=1e-2, rs=None):
    """Compare d/dt fun(arg + t*u) at t=0, for a random unit direction u, with <grad fun, u>.
    Returns True when they agree within rtol/atol."""
    rs = np.random.RandomState() if rs is None else rs
    if isinstance(arg, Value):
        arg = arg.asnumpy()
    direction = rs.standard_normal(np.shape(arg))
    direction /= np.sqrt(np.sum(direction * direction))

    def along(t):
        value = fun(arg + t * direction)
        return value.asnumpy() if isinstance(value, Value) else value
    numeric = (along(eps / 2) - along(-eps / 2)) / eps
    analytic = np.sum(grad(fun)(arg).asnumpy() * direction)
    ok = bool(np.allclose(numeric, analytic, rtol=rtol, atol=atol))
    if verbose:
        print(('Gradient projection OK (numeric grad: {0}, analytic grad: {1})' if ok else
               'Check failed! numeric={0}, analytic={1}').format(numeric, analytic))
    return ok
