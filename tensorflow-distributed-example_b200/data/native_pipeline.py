"""ctypes binding of ``libtfde_runtime.so`` (``csrc/runtime/pipeline.cc``): a background C++ thread runs
shard -> windowed shuffle -> repeat -> batch gather into a prefetch ring; ``__next__`` copies the ready
batch into pinned host tensors (PyTorch's caching host allocator tracks their use by in-flight H2D
copies, so non-blocking copies stay safe). Used automatically by :class:`..dataset.Dataset` for the
``from_tensor_slices -> [shuffle] -> [repeat] -> batch -> [prefetch]`` chains of the reference scripts."""
from __future__ import annotations

import ctypes
import os
import threading
from typing import Optional

import numpy as np
import torch

_LIB_PATH = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "_lib", "libtfde_runtime.so")
_lib = None
_lock = threading.Lock()


def available() -> bool:
    return os.path.exists(_LIB_PATH) and os.environ.get("TFDE_NATIVE_PIPELINE", "1") == "1"


def load():
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                lib = ctypes.CDLL(_LIB_PATH)
                lib.tfde_pipe_create.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p, ctypes.c_int64,
                                                 ctypes.c_int64, ctypes.c_int64, ctypes.c_uint64, ctypes.c_int64,
                                                 ctypes.c_int64, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int]
                lib.tfde_pipe_create.restype = ctypes.c_void_p
                lib.tfde_pipe_next_into.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
                lib.tfde_pipe_next_into.restype = ctypes.c_int64
                lib.tfde_pipe_destroy.argtypes = [ctypes.c_void_p]
                lib.tfde_pipe_destroy.restype = None
                lib.tfde_runtime_abi_version.restype = ctypes.c_int
                _lib = lib
    return _lib


_TORCH_DTYPES = {np.dtype("uint8"): torch.uint8, np.dtype("int8"): torch.int8, np.dtype("int32"): torch.int32,
                 np.dtype("int64"): torch.int64, np.dtype("float32"): torch.float32,
                 np.dtype("float64"): torch.float64, np.dtype("float16"): torch.float16}


class NativeBatchIterator:
    """Iterator over (features, labels) batches (or features only) produced by the C++ runtime."""

    def __init__(self, features: np.ndarray, labels: Optional[np.ndarray], batch: int, shuffle_buffer: int = 0,
                 seed: Optional[int] = None, repeat: int = 1, drop_remainder: bool = False, depth: int = 4,
                 shard_num: int = 1, shard_index: int = 0, pin: Optional[bool] = None):
        self._lib = load()
        self.features = np.ascontiguousarray(features)
        self.labels = None if labels is None else np.ascontiguousarray(labels)
        if self.features.dtype not in _TORCH_DTYPES or (self.labels is not None and self.labels.dtype not in _TORCH_DTYPES):
            raise TypeError("unsupported dtype for the native pipeline")
        self.batch = int(batch)
        n = len(self.features)
        fb = self.features.strides[0] if n else int(np.prod(self.features.shape[1:])) * self.features.itemsize
        lb = 0 if self.labels is None else (self.labels.strides[0] if n else self.labels.itemsize)
        if seed is None:
            seed = int.from_bytes(os.urandom(8), "little")
        self._pin = torch.cuda.is_available() if pin is None else pin
        self._h = self._lib.tfde_pipe_create(self.features.ctypes.data, fb,
                                             None if self.labels is None else self.labels.ctypes.data, lb, n,
                                             int(shuffle_buffer), ctypes.c_uint64(seed & (2 ** 64 - 1)), int(repeat),
                                             self.batch, 1 if drop_remainder else 0, max(2, int(depth)),
                                             int(shard_num), int(shard_index))
        if not self._h:
            raise RuntimeError("tfde_pipe_create failed")

    def __iter__(self):
        return self

    def __next__(self):
        if self._h is None:
            raise StopIteration
        x = torch.empty((self.batch,) + self.features.shape[1:], dtype=_TORCH_DTYPES[self.features.dtype],
                        pin_memory=self._pin)
        y = None
        if self.labels is not None:
            y = torch.empty((self.batch,) + self.labels.shape[1:], dtype=_TORCH_DTYPES[self.labels.dtype],
                            pin_memory=self._pin)
        c = self._lib.tfde_pipe_next_into(self._h, x.data_ptr(), None if y is None else y.data_ptr())
        if c == 0:
            self.close()
            raise StopIteration
        if c < self.batch:
            x = x[:c]
            y = None if y is None else y[:c]
        return x if y is None else (x, y)

    def close(self):
        if self._h is not None:
            self._lib.tfde_pipe_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
