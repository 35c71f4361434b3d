"""Optimizers. The reference uses plain SGD only (distributed_with_keras.py:42 lr=0.001;
mnist_keras_distributed.py:111 ``GradientDescentOptimizer(learning_rate)``; tf2_mnist_distributed.py:137
``SGD(LEARNING_RATE)``); momentum and Adam are the north-star additions of BASELINE.json config 3.

An optimizer here is a *description* (kind + hyper-parameters). The arithmetic runs either as the
epilogue of the fused all-reduce kernel (multi-GPU), as ``tfde_optimizer_apply`` (single GPU), or as
:mod:`..ops.reference` functions on the CPU. Hyper-parameters live in a small device block so a
captured CUDA graph sees learning-rate changes without re-capture.
"""
from __future__ import annotations

import torch

from ..ops import native
from ..ops import reference as ref


class Optimizer:
    kind = "sgd"
    native_code = native.OPT_SGD
    num_slots = 0

    def __init__(self, learning_rate=0.01, name=None):
        self.learning_rate = float(learning_rate)
        self.name = name or type(self).__name__
        self.hparams = None        # device block (CUDA only)
        self.iterations = 0        # host mirror of the applied-step count
        self.momentum = 0.0
        self.beta_1, self.beta_2, self.epsilon = 0.9, 0.999, 1e-7

    # -- device block ---------------------------------------------------------------------------
    def bind(self, device, seed=0, step=0):
        self.iterations = step
        if torch.device(device).type == "cuda":
            self.hparams = native.make_hparams(device, lr=self.learning_rate, momentum=self.momentum,
                                               beta1=self.beta_1, beta2=self.beta_2, eps=self.epsilon,
                                               grad_scale=1.0, seed=seed, step=step)
        return self

    def set_learning_rate(self, lr: float):
        self.learning_rate = float(lr)
        if self.hparams is not None:
            native.hparams_set(self.hparams, lr=lr)

    @property
    def lr(self):
        return self.learning_rate

    # -- single-replica application -----------------------------------------------------------------
    def ensure_slots(self, arena, numel=None):
        n = arena.n_total if numel is None else numel
        if self.num_slots >= 1 and (arena.s1 is None or arena.s1.numel() != n):
            arena.s1 = torch.zeros(n, dtype=torch.float32, device=arena.device)
        if self.num_slots >= 2 and (arena.s2 is None or arena.s2.numel() != n):
            arena.s2 = torch.zeros(n, dtype=torch.float32, device=arena.device)

    def apply_local(self, arena):
        """w <- update(w, g) on the flat arena; moves the (already reduced) metrics tail into w[n:] and
        accumulates it into arena.metrics. Advances the step."""
        self.ensure_slots(arena)
        n = arena.n
        if arena.w.is_cuda:
            native.optimizer_apply(arena.w[:n], arena.g, self.hparams, self.native_code, arena.s1, arena.s2,
                                   arena.w_bf16, n_extra=arena.n_extra, metrics=arena.metrics,
                                   last_extra=arena.w[n:])
        else:
            self._apply_cpu(arena)
            arena.w[n:] = arena.g[n:]
            arena.metrics += arena.g[n:]
        self.iterations += 1

    def _apply_cpu(self, arena):
        n = arena.n
        arena.w[:n] = ref.sgd_update(arena.w[:n], arena.g[:n], self.learning_rate)

    def minimize(self, loss, global_step=None, var_list=None):
        """``optimizer.minimize(loss, global_step)`` inside an Estimator model_fn
        (tf2_mnist_distributed.py:90-91): returns the train op the Estimator runs each step."""
        from ..estimator.symbolic import SymbolicLoss, TrainOp
        if not isinstance(loss, SymbolicLoss):
            raise TypeError("minimize() expects the symbolic loss built inside a model_fn; use "
                            "model.compile()/fit() or train_on_batch() for eager training")
        return TrainOp(loss.model, self, loss, global_step)

    def get_config(self):
        return {"name": self.name, "learning_rate": self.learning_rate}


class SGD(Optimizer):
    """Keras SGD: ``v = momentum*v - lr*g; w += v`` (``w -= lr*g`` when momentum == 0)."""

    def __init__(self, learning_rate=0.01, momentum=0.0, nesterov=False, name=None, lr=None):
        super().__init__(learning_rate if lr is None else lr, name)
        self.momentum = float(momentum)
        self.nesterov = bool(nesterov)
        if self.momentum > 0.0:
            self.kind = "nesterov" if nesterov else "momentum"
            self.native_code = native.OPT_NESTEROV if nesterov else native.OPT_MOMENTUM
            self.num_slots = 1

    def _apply_cpu(self, arena):
        n = arena.n
        if self.momentum == 0.0:
            arena.w[:n] = ref.sgd_update(arena.w[:n], arena.g[:n], self.learning_rate)
        else:
            w, v = ref.momentum_update(arena.w[:n], arena.g[:n], arena.s1[:n], self.learning_rate,
                                       self.momentum, self.nesterov)
            arena.w[:n] = w
            arena.s1[:n] = v

    def get_config(self):
        return {**super().get_config(), "momentum": self.momentum, "nesterov": self.nesterov}


class GradientDescentOptimizer(SGD):
    """tf.train.GradientDescentOptimizer(learning_rate) (mnist_keras_distributed.py:111)."""

    def __init__(self, learning_rate, name="GradientDescent"):
        super().__init__(learning_rate=learning_rate, name=name)


class Adam(Optimizer):
    kind = "adam"
    native_code = native.OPT_ADAM
    num_slots = 2

    def __init__(self, learning_rate=0.001, beta_1=0.9, beta_2=0.999, epsilon=1e-7, name=None):
        super().__init__(learning_rate, name)
        self.beta_1, self.beta_2, self.epsilon = float(beta_1), float(beta_2), float(epsilon)

    def _apply_cpu(self, arena):
        n = arena.n
        w, m, v = ref.adam_update(arena.w[:n], arena.g[:n], arena.s1[:n], arena.s2[:n], self.learning_rate,
                                  self.beta_1, self.beta_2, self.epsilon, self.iterations + 1)
        arena.w[:n] = w
        arena.s1[:n] = m
        arena.s2[:n] = v

    def get_config(self):
        return {**super().get_config(), "beta_1": self.beta_1, "beta_2": self.beta_2, "epsilon": self.epsilon}


def get(identifier):
    if isinstance(identifier, Optimizer):
        return identifier
    if isinstance(identifier, str):
        table = {"sgd": SGD, "adam": Adam}
        if identifier.lower() in table:
            return table[identifier.lower()]()
    raise ValueError(f"unknown optimizer {identifier!r}")
