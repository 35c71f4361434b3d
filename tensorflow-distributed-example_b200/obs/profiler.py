"""ProfilerHook: the (commented-out) ``tf.estimator.ProfilerHook(save_steps=100, output_dir=...,
show_memory=True)`` of mnist_keras_distributed.py:235-237,261. Every ``save_steps`` steps the step is
bracketed with CUDA events (and an NVTX range for external profilers) and a Chrome-trace
``timeline-<step>.json`` is written to ``output_dir`` -- the same artefact TF's hook produced."""
from __future__ import annotations

import json
import os
import time

import torch


class ProfilerHook:
    def __init__(self, save_steps=100, output_dir=".", show_memory=True):
        self.save_steps, self.output_dir, self.show_memory = int(save_steps), output_dir, show_memory
        self._active = None
        self.written = []

    def begin(self, estimator=None):
        os.makedirs(self.output_dir, exist_ok=True)

    def before_run(self, step: int, model=None):
        if step % self.save_steps != 0:
            return
        dev = getattr(model, "device", torch.device("cpu"))
        rec = {"step": step, "t0": time.perf_counter(), "cuda": dev.type == "cuda"}
        if rec["cuda"]:
            stream = getattr(model, "_stream", None) or torch.cuda.current_stream(dev)
            rec["stream"] = stream
            rec["e0"] = torch.cuda.Event(enable_timing=True)
            rec["e1"] = torch.cuda.Event(enable_timing=True)
            rec["e0"].record(stream)
            torch.cuda.nvtx.range_push(f"train_step_{step}")
        self._active = rec

    def after_run(self, step: int, model=None, logs=None):
        rec = self._active
        if rec is None or rec["step"] != step:
            return
        self._active = None
        events = []
        host_us = (time.perf_counter() - rec["t0"]) * 1e6
        events.append({"name": "host: train step", "ph": "X", "pid": 0, "tid": 0, "ts": 0, "dur": host_us})
        if rec["cuda"]:
            rec["e1"].record(rec["stream"])
            torch.cuda.nvtx.range_pop()
            rec["e1"].synchronize()
            dev_us = rec["e0"].elapsed_time(rec["e1"]) * 1e3
            events.append({"name": "device: train step (CUDA events)", "ph": "X", "pid": 1, "tid": 0,
                           "ts": 0, "dur": dev_us})
            if self.show_memory:
                events.append({"name": "memory", "ph": "C", "pid": 1, "ts": 0,
                               "args": {"allocated_MiB": torch.cuda.memory_allocated() / 2 ** 20,
                                        "reserved_MiB": torch.cuda.memory_reserved() / 2 ** 20}})
        path = os.path.join(self.output_dir, f"timeline-{step}.json")
        with open(path, "w") as f:
            json.dump({"traceEvents": events, "displayTimeUnit": "ms"}, f)
        self.written.append(path)

    def end(self, estimator=None):
        pass
