"""Log-level plumbing: ``tf.get_logger().setLevel(args.verbosity)`` (mnist_keras_distributed.py:288-289)
and ``logging.getLogger().setLevel(logging.INFO)`` (tf2_mnist_distributed.py:187)."""
from __future__ import annotations

import logging

_LEVELS = {"DEBUG": logging.DEBUG, "INFO": logging.INFO, "WARN": logging.WARNING, "WARNING": logging.WARNING,
           "ERROR": logging.ERROR, "FATAL": logging.CRITICAL}


def get_logger() -> logging.Logger:
    lg = logging.getLogger("tfde_b200")
    if not lg.handlers:
        h = logging.StreamHandler()
        h.setFormatter(logging.Formatter("%(levelname)s:%(name)s:%(message)s"))
        lg.addHandler(h)
        lg.propagate = False
    return lg


def set_verbosity(level) -> int:
    lv = _LEVELS[level.upper()] if isinstance(level, str) else int(level)
    get_logger().setLevel(lv)
    return lv
