"""TensorBoard event files (chief-only), replacing tf.summary.FileWriter / SummarySaverHook.

Reference: ``save_summary_steps=100`` / ``log_step_count_steps=100`` make the Estimator write ``loss`` and
``global_step/sec`` scalars to ``events.out.tfevents.*`` under ``model_dir``
(mnist_keras_distributed.py:246-247). Uses the installed ``tensorboard`` package's own writer
(``tensorboard.summary.Writer``); if that is unavailable scalars fall back to a JSON-lines file with the
same tags so nothing is silently lost.
"""
from __future__ import annotations

import json
import os
import time


class SummaryWriter:
    def __init__(self, logdir: str):
        self.logdir = logdir
        os.makedirs(logdir, exist_ok=True)
        self._w = None
        self._jsonl = None
        try:
            from tensorboard.summary import Writer
            self._w = Writer(logdir)
        except Exception:
            self._jsonl = open(os.path.join(logdir, "scalars.jsonl"), "a")

    def scalar(self, tag: str, value: float, step: int):
        if self._w is not None:
            self._w.add_scalar(tag, float(value), int(step))
        else:
            self._jsonl.write(json.dumps({"tag": tag, "value": float(value), "step": int(step),
                                          "wall_time": time.time()}) + "\n")

    def flush(self):
        if self._w is not None:
            self._w.flush()
        else:
            self._jsonl.flush()

    def close(self):
        if self._w is not None:
            self._w.close()
        elif self._jsonl is not None:
            self._jsonl.close()


def read_scalars(logdir: str):
    """Reads back every scalar written under ``logdir`` as {tag: [(step, value), ...]} (tests, tools)."""
    out = {}
    jl = os.path.join(logdir, "scalars.jsonl")
    if os.path.exists(jl):
        with open(jl) as f:
            for line in f:
                d = json.loads(line)
                out.setdefault(d["tag"], []).append((d["step"], d["value"]))
        return out
    from tensorboard.backend.event_processing.event_accumulator import EventAccumulator
    acc = EventAccumulator(logdir, size_guidance={"scalars": 0, "tensors": 0})
    acc.Reload()
    tags = acc.Tags()
    for tag in tags.get("scalars", []):
        out[tag] = [(e.step, e.value) for e in acc.Scalars(tag)]
    for tag in tags.get("tensors", []):
        try:
            import numpy as np
            from tensorboard.util import tensor_util
            out[tag] = [(e.step, float(np.asarray(tensor_util.make_ndarray(e.tensor_proto)).reshape(-1)[0]))
                        for e in acc.Tensors(tag)]
        except Exception:
            pass
    return out
