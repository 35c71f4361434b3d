"""ProfilerHook: the (commented-out) ``tf.estimator.ProfilerHook(save_steps=100, output_dir=...,
show_memory=True)`` of mnist_keras_distributed.py:235-237,261. Every ``save_steps`` steps the step is
bracketed with CUDA events (and an NVTX range for external profilers) and a Chrome-trace
``timeline-<step>.json`` is written to ``output_dir`` -- the same artefact TF's hook produced: the sampled step runs
span by span (:class:`StepTracer`), so the trace has one device span per layer and direction, per loss / reduction /
optimizer kernel, each naming the native entry points it launched."""
from __future__ import annotations

import json
import os
import time

import torch


import contextlib


class StepTracer:
    """Per-kernel timeline of ONE training step (what TF's ProfilerHook showed per op): every layer's forward and
    backward, the loss, the cross-replica reduction and the optimizer are bracketed by a CUDA-event pair on the
    model's stream (and an NVTX range, for external profilers); each span lists the native entry points
    (``tfde_*`` kernels) launched inside it. A step that runs as the single whole-step kernel is resolved further
    with the kernel's own in-kernel ``globaltimer`` stamps (one span per phase)."""

    def __init__(self, device):
        self.device = torch.device(device)
        self.cuda = self.device.type == "cuda"
        self.spans = []          # (name, e0, e1, host_t0, host_t1, [entry points])
        self.extra = []          # ready-made trace events (in-kernel phases)
        self.t_origin = time.perf_counter()

    @contextlib.contextmanager
    def span(self, name: str):
        from ..ops import native
        sink = []
        prev = native.set_trace_sink(sink) if self.cuda else None
        e0 = e1 = None
        if self.cuda:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.nvtx.range_push(name)
            e0.record(torch.cuda.current_stream(self.device))
        h0 = time.perf_counter()
        try:
            yield
        finally:
            h1 = time.perf_counter()
            if self.cuda:
                e1.record(torch.cuda.current_stream(self.device))
                torch.cuda.nvtx.range_pop()
                native.set_trace_sink(prev)
            self.spans.append((name, e0, e1, h0, h1, sink))

    def add_kernel_phases(self, kernel: str, phases):
        """phases: [(name, start_us, dur_us)] relative to the kernel's first stamp."""
        self._phases = (kernel, list(phases))

    def events(self):
        out = []
        if self.cuda and self.spans:
            torch.cuda.synchronize(self.device)
            first = self.spans[0][1]
        kernel_span_ts = None
        for name, e0, e1, h0, h1, sink in self.spans:
            args = {"kernels": list(sink)} if sink else {}
            out.append({"name": "host: " + name, "ph": "X", "pid": 0, "tid": 0, "ts": (h0 - self.t_origin) * 1e6,
                        "dur": (h1 - h0) * 1e6, "args": args})
            if self.cuda:
                ts = first.elapsed_time(e0) * 1e3
                dur = e0.elapsed_time(e1) * 1e3
                out.append({"name": name, "ph": "X", "pid": 1, "tid": 0, "ts": ts, "dur": dur, "args": args})
                if getattr(self, "_phases", None) and sink and any(self._phases[0] in k for k in sink):
                    kernel_span_ts = ts
        if getattr(self, "_phases", None):
            kernel, phases = self._phases
            base = kernel_span_ts or 0.0
            for name, start, dur in phases:
                out.append({"name": f"{kernel}: {name}", "ph": "X", "pid": 1, "tid": 1, "ts": base + start, "dur": dur})
        return out


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
        if model is not None and hasattr(model, "start_trace"):
            rec["tracer"] = model.start_trace()      # this step runs span by span (no CUDA graph, no fast path)
        self._active = rec

    def after_run(self, step: int, model=None, logs=None):
        rec = self._active
        if rec is None or rec["step"] != step:
            return
        self._active = None
        events = []
        tracer = rec.get("tracer")
        if tracer is not None:
            model.stop_trace()
            events += tracer.events()
        host_us = (time.perf_counter() - rec["t0"]) * 1e6
        events.append({"name": "host: train step", "ph": "X", "pid": 0, "tid": 2, "ts": 0, "dur": host_us})
        if rec["cuda"]:
            rec["e1"].record(rec["stream"])
            torch.cuda.nvtx.range_pop()
            rec["e1"].synchronize()
            dev_us = rec["e0"].elapsed_time(rec["e1"]) * 1e3
            events.append({"name": "device: train step (CUDA events)", "ph": "X", "pid": 1, "tid": 2,
                           "ts": 0, "dur": dev_us})
            if self.show_memory:
                events.append({"name": "memory", "ph": "C", "pid": 1, "ts": 0,
                               "args": {"allocated_MiB": torch.cuda.memory_allocated() / 2 ** 20,
                                        "reserved_MiB": torch.cuda.memory_reserved() / 2 ** 20}})
        os.makedirs(self.output_dir, exist_ok=True)
        path = os.path.join(self.output_dir, f"timeline-{step}.json")
        with open(path, "w") as f:
            json.dump({"traceEvents": events, "displayTimeUnit": "ms"}, f)
        self.written.append(path)

    def end(self, estimator=None):
        pass
