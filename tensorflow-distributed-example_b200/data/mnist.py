"""MNIST sources.

The reference loads MNIST two ways: ``tfds.load('mnist', data_dir='/tmp/data', with_info=True,
as_supervised=True)`` (distributed_with_keras.py:25-28) and ``tf.keras.datasets.mnist.load_data()``
(mnist_keras_distributed.py:207-208). Both download from the network; this box has none, so:

1. if a real ``mnist.npz`` (Keras layout: x_train, y_train, x_test, y_test) is found in
   ``$TFDE_MNIST_PATH``, ``<data_dir>/mnist.npz`` or ``~/.keras/datasets/mnist.npz`` it is used;
2. otherwise a deterministic *synthetic* MNIST-shaped set is generated: uint8 ``[N,28,28]`` images whose
   pixels depend on the label (a per-class stroke template + noise), so a model can actually learn it
   and loss curves are meaningful. Benchmarks say ``"data": "synthetic"``.
"""
from __future__ import annotations

import os
from typing import Dict, Optional, Tuple

import numpy as np

from .dataset import Dataset

TRAIN_N, TEST_N = 60000, 10000


def _find_npz(data_dir: Optional[str]) -> Optional[str]:
    cands = [os.environ.get("TFDE_MNIST_PATH")]
    if data_dir:
        cands.append(os.path.join(data_dir, "mnist.npz"))
    cands.append(os.path.expanduser("~/.keras/datasets/mnist.npz"))
    for c in cands:
        if c and os.path.isfile(c):
            return c
    return None


def _templates(rng: np.random.Generator) -> np.ndarray:
    """10 smooth 28x28 class templates in [0,1]."""
    t = np.zeros((10, 28, 28), dtype=np.float32)
    yy, xx = np.mgrid[0:28, 0:28].astype(np.float32)
    for c in range(10):
        for _ in range(3):
            cy, cx = rng.uniform(6, 22, size=2)
            ang = rng.uniform(0, np.pi)
            d = np.abs((yy - cy) * np.cos(ang) - (xx - cx) * np.sin(ang))
            along = np.abs((yy - cy) * np.sin(ang) + (xx - cx) * np.cos(ang))
            t[c] += np.exp(-(d ** 2) / 3.0) * (along < rng.uniform(5, 10))
    return np.clip(t, 0, 1)


def synthetic_mnist(n_train: int = TRAIN_N, n_test: int = TEST_N, seed: int = 1234):
    rng = np.random.default_rng(seed)
    tmpl = _templates(rng)

    def make(n):
        y = rng.integers(0, 10, size=n, dtype=np.int64)
        noise = rng.normal(0.0, 0.25, size=(n, 28, 28)).astype(np.float32)
        shift = rng.integers(-2, 3, size=(n, 2))
        x = tmpl[y]
        # cheap per-example jitter: roll by a small random offset along each axis group-wise
        for dy in range(-2, 3):
            for dx in range(-2, 3):
                sel = (shift[:, 0] == dy) & (shift[:, 1] == dx)
                if sel.any():
                    x[sel] = np.roll(np.roll(tmpl[y[sel]], dy, axis=1), dx, axis=2)
        img = np.clip((x + noise) * 255.0, 0, 255).astype(np.uint8)
        return img, y.astype(np.uint8)

    return make(n_train), make(n_test)


_cache: Dict[tuple, tuple] = {}


def load_data(data_dir: Optional[str] = None, n_train: int = TRAIN_N, n_test: int = TEST_N):
    """``tf.keras.datasets.mnist.load_data()``: ((x_train, y_train), (x_test, y_test)), uint8."""
    path = _find_npz(data_dir)
    key = (path, n_train, n_test)
    if key in _cache:
        return _cache[key]
    if path:
        with np.load(path) as f:
            out = ((f["x_train"][:n_train], f["y_train"][:n_train]), (f["x_test"][:n_test], f["y_test"][:n_test]))
    else:
        out = synthetic_mnist(n_train, n_test)
    _cache[key] = out
    return out


def is_synthetic(data_dir: Optional[str] = None) -> bool:
    return _find_npz(data_dir) is None


class DatasetInfo:
    def __init__(self, name, splits, synthetic):
        self.name, self.splits, self.synthetic = name, splits, synthetic
        self.features = {"image": (28, 28, 1), "label": 10}

    def __repr__(self):
        return f"DatasetInfo(name={self.name!r}, splits={self.splits}, synthetic={self.synthetic})"


def load(name="mnist", data_dir=None, with_info=False, as_supervised=True, n_train=TRAIN_N, n_test=TEST_N):
    """``tfds.load`` analogue: {'train': Dataset, 'test': Dataset} of (image uint8 [28,28,1], label int64)."""
    if name != "mnist":
        raise ValueError("only 'mnist' is available")
    (xtr, ytr), (xte, yte) = load_data(data_dir, n_train, n_test)
    ds = {"train": Dataset.from_tensor_slices((xtr[..., None], ytr.astype(np.int64))),
          "test": Dataset.from_tensor_slices((xte[..., None], yte.astype(np.int64)))}
    if not as_supervised:
        raise ValueError("as_supervised=False (dict elements) is not supported")
    info = DatasetInfo(name, {"train": len(xtr), "test": len(xte)}, is_synthetic(data_dir))
    return (ds, info) if with_info else ds


def disable_progress_bar():
    """tfds.disable_progress_bar() (distributed_with_keras.py:10): nothing to disable here."""
