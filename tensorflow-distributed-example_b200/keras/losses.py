"""Losses. The reference uses exactly one loss in three spellings:

* ``SparseCategoricalCrossentropy(from_logits=True)`` (distributed_with_keras.py:41),
* the same with ``reduction=NONE`` + manual ``reduce_sum(loss) * (1/BATCH_SIZE)``
  (tf2_mnist_distributed.py:81-83),
* the string ``'sparse_categorical_crossentropy'`` applied to softmax probabilities
  (mnist_keras_distributed.py:108,114).

Here the loss object only carries configuration; the fused kernel ``tfde_xent`` computes the loss sum,
the correct-prediction count and the gradient in one pass, already multiplied by ``1/global_batch`` so
that SUM-reduced gradients equal the global-batch mean gradient (S5).
"""
from __future__ import annotations


class Reduction:
    AUTO = "auto"
    NONE = "none"
    SUM = "sum"
    SUM_OVER_BATCH_SIZE = "sum_over_batch_size"


class Loss:
    from_probs = False
    name = "loss"


class SparseCategoricalCrossentropy(Loss):
    def __init__(self, from_logits=False, reduction=Reduction.AUTO, name="sparse_categorical_crossentropy"):
        self.from_logits = bool(from_logits)
        self.reduction = reduction
        self.name = name

    @property
    def from_probs(self):
        return not self.from_logits

    def __call__(self, y_true, y_pred):
        """Eager evaluation (custom-loop idiom of tf2_mnist_distributed.py:81-83). Honors ``reduction``:
        NONE returns the per-example vector."""
        import torch
        from ..ops import reference as ref
        from ..estimator.symbolic import SymbolicLoss, SymbolicOutput
        if isinstance(y_pred, SymbolicOutput):    # model_fn graph mode: defer to the fused loss kernel
            return SymbolicLoss(y_pred.model, self, reduced=self.reduction != Reduction.NONE)
        y_pred = y_pred.detach().float().cpu()
        labels = y_true.reshape(-1).long().cpu()
        if self.from_logits:
            per = torch.logsumexp(y_pred, dim=-1) - y_pred.gather(1, labels[:, None]).squeeze(1)
        else:
            per = -torch.log(y_pred.gather(1, labels[:, None]).squeeze(1).clamp(ref.EPS_PROB, 1 - ref.EPS_PROB))
        if self.reduction == Reduction.NONE:
            return per
        if self.reduction == Reduction.SUM:
            return per.sum()
        return per.mean()

    def get_config(self):
        return {"name": self.name, "from_logits": self.from_logits, "reduction": self.reduction}


def get(identifier) -> Loss:
    if isinstance(identifier, Loss):
        return identifier
    if isinstance(identifier, str) and identifier.lower() in ("sparse_categorical_crossentropy", "scce"):
        return SparseCategoricalCrossentropy(from_logits=False)
    raise ValueError(f"unsupported loss {identifier!r}; the reference uses sparse categorical cross-entropy only")
