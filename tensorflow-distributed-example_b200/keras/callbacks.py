"""Callbacks for ``fit`` (the Keras half of the "hook list on the train spec" mechanism).
``History`` is what ``fit`` returns (distributed_with_keras.py:63, T1)."""
from __future__ import annotations

import time
from typing import Dict, List


class Callback:
    model = None

    def set_model(self, model):
        self.model = model

    def on_train_begin(self, logs=None): ...
    def on_train_end(self, logs=None): ...
    def on_epoch_begin(self, epoch, logs=None): ...
    def on_epoch_end(self, epoch, logs=None): ...
    def on_train_batch_begin(self, batch, logs=None): ...
    def on_train_batch_end(self, batch, logs=None): ...


class History(Callback):
    def __init__(self):
        self.history: Dict[str, List[float]] = {}
        self.epoch: List[int] = []

    def on_epoch_end(self, epoch, logs=None):
        self.epoch.append(epoch)
        for k, v in (logs or {}).items():
            self.history.setdefault(k, []).append(v)


class CallbackList:
    def __init__(self, callbacks, model):
        self.callbacks = list(callbacks)
        for c in self.callbacks:
            c.set_model(model)
        # skip the per-batch python dispatch entirely when nobody overrides it (hot loop)
        self._batch_begin = [c for c in self.callbacks
                             if type(c).on_train_batch_begin is not Callback.on_train_batch_begin]
        self._batch_end = [c for c in self.callbacks
                           if type(c).on_train_batch_end is not Callback.on_train_batch_end]

    def on_train_begin(self):
        for c in self.callbacks:
            c.on_train_begin()

    def on_train_end(self):
        for c in self.callbacks:
            c.on_train_end()

    def on_epoch_begin(self, epoch):
        for c in self.callbacks:
            c.on_epoch_begin(epoch)

    def on_epoch_end(self, epoch, logs):
        for c in self.callbacks:
            c.on_epoch_end(epoch, logs)

    def on_train_batch_begin(self, batch):
        for c in self._batch_begin:
            c.on_train_batch_begin(batch)

    def on_train_batch_end(self, batch):
        for c in self._batch_end:
            c.on_train_batch_end(batch)


class ModelCheckpoint(Callback):
    """Chief-only weight checkpoint every ``save_freq`` epochs."""

    def __init__(self, filepath, save_freq=1):
        self.filepath, self.save_freq = filepath, save_freq

    def on_epoch_end(self, epoch, logs=None):
        if (epoch + 1) % self.save_freq == 0 and self.model.strategy.is_chief:
            self.model.save_weights(self.filepath.format(epoch=epoch + 1))


class TensorBoard(Callback):
    """Chief-only scalar summaries per epoch (loss / accuracy / images per second)."""

    def __init__(self, log_dir):
        self.log_dir = log_dir
        self._w = None
        self._t = None

    def on_train_begin(self, logs=None):
        if self.model.strategy.is_chief:
            from ..obs.summary import SummaryWriter
            self._w = SummaryWriter(self.log_dir)

    def on_epoch_begin(self, epoch, logs=None):
        self._t = time.time()

    def on_epoch_end(self, epoch, logs=None):
        if self._w is not None:
            for k, v in (logs or {}).items():
                self._w.scalar(f"epoch_{k}", v, epoch)
            self._w.flush()

    def on_train_end(self, logs=None):
        if self._w is not None:
            self._w.close()


class LearningRateScheduler(Callback):
    """schedule(epoch, lr) -> lr. Works with captured CUDA graphs: the rate lives in device memory."""

    def __init__(self, schedule):
        self.schedule = schedule

    def on_epoch_begin(self, epoch, logs=None):
        opt = self.model.optimizer
        opt.set_learning_rate(self.schedule(epoch, opt.learning_rate))
