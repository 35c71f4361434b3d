"""Registry that routes the filter-gradient GEMM of selected tensors to ``tfde_dense_wgrad_push`` (the
GEMM whose epilogue also stores its tiles into the peers' all-reduce receive slots)."""
from __future__ import annotations

import ctypes

_registry = {}
stats = {"pushed_launches": 0}


def register(grad_ptr: int, desc) -> None:
    _registry[int(grad_ptr)] = desc


def lookup(grad_ptr: int):
    return _registry.get(int(grad_ptr))


def clear() -> None:
    _registry.clear()
