"""Logger factory used across the front end (role of the reference's minpy/utils/log.py:65-80).

Loggers default to WARNING so that the per-primitive debug messages cost one level check.
"""
from __future__ import absolute_import

import logging
import os
import sys

_LEVEL = getattr(logging, os.environ.get('MINPY_LOG_LEVEL', 'WARNING').upper(), logging.WARNING)
_FORMAT = '%(levelname).1s %(asctime)s %(name)s:%(lineno)d] %(message)s'
_configured = set()


def get_logger(name=None, level=None):
    """Return a logger that writes one formatted line per record to stderr."""
    logger = logging.getLogger(name)
    if name not in _configured:
        handler = logging.StreamHandler(sys.stderr)
        handler.setFormatter(logging.Formatter(_FORMAT, datefmt='%m%d %H:%M:%S'))
        logger.addHandler(handler)
        logger.propagate = False
        _configured.add(name)
    logger.setLevel(_LEVEL if level is None else level)
    return logger


def set_level(level):
    """Change the level of every logger handed out so far and of future ones."""
    global _LEVEL
    _LEVEL = level
    for name in _configured:
        logging.getLogger(name).setLevel(level)
