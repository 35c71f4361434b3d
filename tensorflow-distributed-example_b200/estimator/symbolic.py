"""Graph-mode shim for hand-written Estimator ``model_fn``s (capability 11).

TF1 Estimators call ``model_fn(features, labels, mode)`` ONCE to build a graph. The reference's
``model_fn`` (tf2_mnist_distributed.py:65-91) builds a Keras model, takes the per-example loss with
``reduction=NONE``, scales it by hand -- ``tf.reduce_sum(loss) * (1. / BATCH_SIZE)`` -- and returns
``optimizer.minimize(loss, global_step)``. There is no autograd graph in this framework (backward is
explicit kernels), so the same code is expressed through tiny symbolic handles: the model call returns a
:class:`SymbolicOutput`, the loss call a :class:`SymbolicLoss` that remembers reduction and scale, and
``minimize`` a :class:`TrainOp` the Estimator executes with the fused step
(``Sequential.train_on_batch(global_batch = 1/scale)``).
"""
from __future__ import annotations

from typing import Optional


class SymbolicTensor:
    def __init__(self, name, shape=None, dtype="float32"):
        self.name, self.shape, self.dtype = name, shape, dtype

    def __repr__(self):
        return f"<SymbolicTensor {self.name} shape={self.shape} dtype={self.dtype}>"


def placeholder(dtype="float32", shape=None, name="Placeholder") -> SymbolicTensor:
    """tf.placeholder(tf.float32, [None, 784]) (mnist_keras_distributed.py:159)."""
    return SymbolicTensor(name, list(shape) if shape is not None else None, str(dtype))


class SymbolicOutput(SymbolicTensor):
    def __init__(self, model, inputs, training):
        super().__init__(f"{model.name}/output")
        self.model, self.inputs, self.training = model, inputs, training


class SymbolicLoss:
    def __init__(self, model, loss_obj, reduced=False, scale: Optional[float] = None):
        self.model, self.loss_obj, self.reduced, self.scale = model, loss_obj, reduced, scale

    def _scaled(self, s):
        return SymbolicLoss(self.model, self.loss_obj, self.reduced, (self.scale or 1.0) * float(s))

    __mul__ = __rmul__ = lambda self, s: self._scaled(s)

    def __truediv__(self, s):
        return self._scaled(1.0 / float(s))

    @property
    def global_batch(self) -> Optional[int]:
        """``reduce_sum(loss) * (1/GLOBAL_BATCH)`` -> the global batch the gradient is normalised by."""
        if self.reduced and self.scale:
            return int(round(1.0 / self.scale))
        return None


def reduce_sum(x):
    if isinstance(x, SymbolicLoss):
        return SymbolicLoss(x.model, x.loss_obj, True, x.scale)
    import torch
    return torch.as_tensor(x).sum()


def reduce_mean(x):
    if isinstance(x, SymbolicLoss):
        return SymbolicLoss(x.model, x.loss_obj, True, None)   # mean over the per-replica batch
    import torch
    return torch.as_tensor(x).float().mean()


class TrainOp:
    def __init__(self, model, optimizer, loss: SymbolicLoss, global_step=None):
        self.model, self.optimizer, self.loss, self.global_step = model, optimizer, loss, global_step


class GlobalStep:
    def __init__(self):
        self.value = 0


_global_step = GlobalStep()


def get_or_create_global_step() -> GlobalStep:
    """tf.compat.v1.train.get_or_create_global_step() (tf2_mnist_distributed.py:91)."""
    return _global_step
