"""`mxnet.symbol` -> minpy_b200.symbol (see mxnet/__init__.py)."""
from minpy_b200.symbol import *  # noqa: F401,F403
from minpy_b200.symbol import Symbol, Variable  # noqa: F401
