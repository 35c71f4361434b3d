"""Keras-like ``Sequential`` with ``compile / fit / evaluate / predict / summary`` and weight I/O.

Reference surface: ``tf.keras.Sequential([...])`` + ``compile(loss, optimizer, metrics)`` +
``fit(x, epochs, steps_per_epoch)`` (distributed_with_keras.py:33-43,63), ``model.summary()``
(mnist_keras_distributed.py:117).

Execution model (B200-first, no tracing compiler): ``build()`` lays every trainable variable into ONE
flat fp32 parameter arena and every gradient into ONE flat gradient arena (what the fused
all-reduce+optimizer kernel consumes in a single launch); a fusion pass merges Conv+ReLU+MaxPool and
BN+ReLU into single kernels; the training step is a fixed sequence of hand-written kernels with static
buffers, so after two eager warm-up steps the whole step (forward, loss, backward, cross-replica
reduction, update, metric accumulation) is captured once in a CUDA graph and replayed with one launch.
"""
from __future__ import annotations

import contextlib
import gc
import logging
import os
import time
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from .. import ops
from ..parallel.comm import Arena
from ..parallel.strategy import get_strategy
from . import layers as L
from . import losses as losses_mod
from . import optimizers as opt_mod
from .callbacks import CallbackList, History

log = logging.getLogger("tfde_b200")

_global_seed: Optional[int] = None


def set_seed(seed: int):
    """tf.random.set_seed analogue: makes initialisation and dropout reproducible."""
    global _global_seed
    _global_seed = int(seed)


def _fresh_seed() -> int:
    if _global_seed is not None:
        return _global_seed
    return int.from_bytes(os.urandom(4), "little")


@contextlib.contextmanager
def _capture(graph, stream):
    """torch.cuda.graph with the cyclic garbage collector paused. A collection that happens to run in
    the middle of a capture can finalise CUDA graphs of dead models (CUDAGraph.reset frees device memory),
    which CUDA forbids while a stream is capturing and which invalidates the capture in progress."""
    gc.collect()
    was_enabled = gc.isenabled()
    gc.disable()
    try:
        with torch.cuda.graph(graph, stream=stream):
            yield
    finally:
        if was_enabled:
            gc.enable()


class _Slot:
    """Pinned host staging for one in-flight step (inputs in, metrics tail out) + its captured graph."""

    def __init__(self, x_shape, x_dtype, batch):
        self.host_x = torch.empty(x_shape, dtype=x_dtype).pin_memory()
        self.host_y = torch.empty((batch,), dtype=torch.int64).pin_memory()
        self.host_tail = torch.zeros(4, dtype=torch.float32).pin_memory()
        self.graph = None
        self.event = None
        self.dev_y = None      # CUDA-tensor aliases of the pinned buffers (zero-copy), created lazily
        self.dev_tail = None
        self.overlap = False   # image H2D on a copy stream into dev_x, overlapping the previous step
        self.dev_x = None
        self.copy_event = None


class _GraphState:
    def __init__(self):
        self.in_x = None
        self.in_y = None
        self.eager_steps = 0
        self.graph = None      # step only (device-resident inputs)
        self.slots = None      # two _Slot: step + H2D/D2H staging
        self.turn = 0


class Sequential:
    def __init__(self, layers: Optional[Sequence[L.Layer]] = None, name: Optional[str] = None,
                 seed: Optional[int] = None):
        self.name = name or "sequential"
        self.strategy = get_strategy()
        self.device = self.strategy.device
        self.replica_id = self.strategy.replica_id
        self.layers: List[L.Layer] = []
        self.built = False
        self.compiled = False
        self.seed = _fresh_seed() if seed is None else int(seed)
        self.arena: Optional[Arena] = None
        self.state: Optional[torch.Tensor] = None     # non-trainable flat buffer (BN moving stats)
        self.variables: List[L.Variable] = []
        self.exec_plan: List[L.Layer] = []
        self.loss = None
        self.optimizer: Optional[opt_mod.Optimizer] = None
        self.metrics_names: List[str] = ["loss"]
        self.use_cuda_graph = os.environ.get("TFDE_CUDA_GRAPH", "1") == "1"
        self.graph_warmup_steps = 2
        self._graphs: Dict[tuple, _GraphState] = {}
        self._stream = None
        self._type_counts: Dict[str, int] = {}
        self.stop_training = False
        self.gpu_launches_per_step = 0
        self.steps_per_execution = 1
        self._fused = None           # ops.fused_step.FusedStepPlan once the whole-step kernel is in use
        self._fused_checked = False
        self._tracer = None          # obs.profiler.StepTracer while a ProfilerHook samples this step
        for l in layers or []:
            self.add(l)

    # ------------------------------------------------------------------------------------------
    # construction
    # ------------------------------------------------------------------------------------------
    def add(self, layer: L.Layer):
        if self.built:
            raise RuntimeError("cannot add layers after the model was built")
        if layer.name is None:
            k = self._type_counts.get(layer._type_name, 0)
            layer.name = layer._type_name if k == 0 else f"{layer._type_name}_{k}"
            self._type_counts[layer._type_name] = k + 1
        layer.model = self
        if isinstance(layer, L.Dropout):
            layer.stream_id = len(self.layers) + 1   # RNG stream = position in the model (deterministic)
        self.layers.append(layer)
        if len(self.layers) == 1 and layer.input_shape_arg is not None:
            pass  # built lazily in build()/compile()
        return self

    @property
    def hparams(self):
        return self.optimizer.hparams if self.optimizer is not None else None

    @property
    def host_step(self) -> int:
        return self.optimizer.iterations if self.optimizer is not None else 0

    def build(self, input_shape=None):
        """``input_shape`` excludes the batch dimension."""
        if self.built:
            return
        if input_shape is None:
            input_shape = self.layers[0].input_shape_arg if self.layers else None
        if input_shape is None:
            raise ValueError("the first layer needs input_shape=... (or call build(input_shape))")
        shape = tuple(int(s) for s in input_shape)
        self.input_shape = shape
        first_weighted = True
        for layer in self.layers:
            shape = layer.build(shape)
            if layer.variables and first_weighted:
                layer.need_dx = False   # nothing upstream needs a data gradient
                first_weighted = False
            elif not first_weighted:
                layer.need_dx = True
        self.output_shape = shape
        self.variables = [v for layer in self.layers for v in layer.variables]
        self._allocate()
        self._initialize()
        self._fuse()
        self.built = True

    def _allocate(self):
        def lay(vs):
            off = 0
            for v in vs:
                v.offset = off
                off += (v.size + 7) // 8 * 8   # slices 32-byte aligned in fp32, 16-byte aligned in the bf16 shadow (TMA)
            return off

        train = [v for v in self.variables if v.trainable]
        frozen = [v for v in self.variables if not v.trainable]
        n = lay(train)
        m = lay(frozen)
        comm = self.strategy.comm() if self.strategy.num_replicas_in_sync > 1 else None
        self.arena = Arena(n, self.device, comm=comm)
        self.state = torch.zeros(max(m, 4), dtype=torch.float32, device=self.device)
        for v in train:
            v.value = self.arena.w[v.offset:v.offset + v.size].view(v.shape)
            v.grad = self.arena.g[v.offset:v.offset + v.size].view(v.shape)
        for v in frozen:
            v.value = self.state[v.offset:v.offset + v.size].view(v.shape)
        # bf16 activations: the tensor-core GEMMs read the weights from ONE bf16 shadow of the parameter arena,
        # refreshed by a single cast kernel per step (instead of one cast/transpose kernel per layer and GEMM)
        from ..ops import mixed
        self._w16 = None
        if self.device.type == "cuda" and mixed.activation_dtype() == "bfloat16":
            self._w16 = torch.zeros(self.arena.n_total, dtype=torch.bfloat16, device=self.device)
            for v in train:
                v.value16 = self._w16[v.offset:v.offset + v.size].view(v.shape)

    def _refresh_w16(self):
        if getattr(self, "_w16", None) is not None:
            from ..ops import mixed
            mixed.cast_bf16_masked(self.arena.w, None, out=self._w16)

    def _initialize(self):
        gen = torch.Generator(device="cpu")
        gen.manual_seed(self.seed & 0x7FFFFFFFFFFFFFFF)
        self.arena.w.zero_()
        for v in self.variables:
            v.value.copy_(v.initializer(v.shape, gen).to(self.device))
        if self.strategy.num_replicas_in_sync > 1:
            # initial values come from the chief (N3 / G13); also aligns the seed used by dropout
            self.strategy.broadcast_(self.arena.w, src=0)
            self.strategy.broadcast_(self.state, src=0)

    def _fuse(self):
        """Fusion pass: Conv2D(relu)+MaxPool(2x2) -> one kernel (first layer of Model A);
        BatchNormalization + Activation('relu') -> BN with fused ReLU (Model B)."""
        plan: List[L.Layer] = []
        i = 0
        ls = self.layers
        allow = os.environ.get("TFDE_FUSION", "1") == "1"
        while i < len(ls):
            cur = ls[i]
            nxt = ls[i + 1] if i + 1 < len(ls) else None
            if allow and nxt is not None and L.FusedConvReluPool.matches(cur, nxt):
                plan.append(L.FusedConvReluPool(cur, nxt))
                i += 2
                continue
            if (allow and isinstance(cur, L.BatchNormalization) and isinstance(nxt, L.Activation)
                    and nxt.activation == "relu"):
                cur.fused_relu = True
                nxt.fused_away = True
            if isinstance(cur, L.Conv2D) and cur.activation is None and isinstance(nxt, L.BatchNormalization):
                # d(loss)/d(bias) of a convolution feeding a training-mode batch-norm is identically zero
                # (batch-norm removes any per-channel shift); the bf16-activation path skips that reduction
                cur.bias_grad_zero = True
                cur.feeds_bn, nxt.stats_source = True, cur     # BN statistics come out of the conv GEMM's epilogue
            plan.append(cur)
            i += 1
        self.exec_plan = plan

    # ------------------------------------------------------------------------------------------
    # compile
    # ------------------------------------------------------------------------------------------
    def compile(self, optimizer="sgd", loss=None, metrics=None, steps_per_execution=1, **_ignored):
        """``steps_per_execution`` (Keras): how many optimizer steps ``fit`` hands to the device per launch. The
        whole-step kernel runs them back to back inside one launch, each step with its own input batch."""
        self.steps_per_execution = max(1, int(steps_per_execution or 1))
        self.loss = losses_mod.get(loss)
        self.optimizer = opt_mod.get(optimizer)
        self.metrics_names = ["loss"]
        for m in metrics or []:
            if m in ("accuracy", "acc", "sparse_categorical_accuracy"):
                self.metrics_names.append("accuracy")
            else:
                raise ValueError(f"unsupported metric {m!r}; the reference uses 'accuracy' only")
        self.compiled = True
        if self.layers and self.layers[0].input_shape_arg is not None and not self.built:
            self.build()
        if self.built:
            self._bind_optimizer()

    def _bind_optimizer(self):
        if getattr(self.optimizer, "_bound_to", None) is self:
            return
        self.optimizer.bind(self.device, seed=self.seed, step=self.optimizer.iterations)
        self.optimizer._bound_to = self
        if self.strategy.num_replicas_in_sync > 1:
            ch = self.strategy.comm().prepare(self.arena, self.optimizer)
            if hasattr(ch, "enable_fused_push"):
                # the largest Dense kernel gradient is pushed to the peers by its own wgrad GEMM
                dense = [l for l in self.exec_plan if isinstance(l, L.Dense) and l.kernel.size % 4 == 0]
                if dense:
                    big = max(dense, key=lambda l: l.kernel.size)
                    if big.kernel.size * 4 >= (256 << 10):
                        ch.enable_fused_push(big.kernel.grad, big.kernel.offset)
        else:
            self.optimizer.ensure_slots(self.arena)

    # ------------------------------------------------------------------------------------------
    # forward / backward
    # ------------------------------------------------------------------------------------------
    def _prep_input(self, x: torch.Tensor) -> torch.Tensor:
        """Accepts [B,H,W], [B,H,W,C] or flat [B,prod] inputs (uint8 or float)."""
        want = (x.shape[0],) + self.input_shape
        if tuple(x.shape) != want:
            n = 1
            for s in self.input_shape:
                n *= s
            if x[0].numel() != n:
                raise ValueError(f"input of shape {tuple(x.shape)} does not match model input {self.input_shape}")
            x = x.reshape(want)
        if x.dtype == torch.uint8:
            first = self.exec_plan[0]
            if not isinstance(first, L.FusedConvReluPool):
                x = ops.u8_to_f32(x, 1.0 / 255.0, out=self._in_buf("x_f32", x.shape))
        elif x.dtype != torch.float32:
            x = x.to(torch.float32)
        return x

    def _in_buf(self, tag, shape, dtype=torch.float32):
        key = ("__model__", tag, tuple(shape), dtype)
        b = self._graphs.get(key)
        if b is None:
            b = torch.empty(tuple(shape), dtype=dtype, device=self.device)
            self._graphs[key] = b
        return b

    def _span(self, name: str):
        return self._tracer.span(name) if self._tracer is not None else contextlib.nullcontext()

    def start_trace(self):
        """The next training step runs span by span (eagerly, layer kernels bracketed by CUDA events): used by
        ``ProfilerHook`` on its sampled steps. Returns the tracer; ``stop_trace()`` ends the mode."""
        from ..obs.profiler import StepTracer
        self._tracer = StepTracer(self.device)
        return self._tracer

    def stop_trace(self):
        self._tracer = None

    def _forward(self, x, training: bool):
        self._refresh_w16()
        x = self._prep_input(x)
        for node in self.exec_plan:
            with self._span(f"{type(node).__name__}/{node.name}/forward"):
                x = node.run(x, training)
        return x

    def _backward(self, dy):
        for node in reversed(self.exec_plan):
            if dy is None:
                break
            with self._span(f"{type(node).__name__}/{node.name}/backward"):
                dy = node.run_back(dy)

    def __call__(self, x, training=False):
        from ..estimator.symbolic import SymbolicOutput, SymbolicTensor
        if isinstance(x, SymbolicTensor):     # graph-building call inside an Estimator model_fn
            return SymbolicOutput(self, x, training)
        if not self.built:
            self.build(tuple(x.shape[1:]))
        x = self._to_device(x)
        return self._forward(x, training)

    def _to_device(self, t, dtype=None):
        if isinstance(t, np.ndarray):
            t = torch.from_numpy(np.ascontiguousarray(t))
        if dtype is not None and t.dtype != dtype:
            t = t.to(dtype)
        if t.device != self.device:
            t = t.to(self.device, non_blocking=True)
        return t.contiguous()

    # ------------------------------------------------------------------------------------------
    # the training step
    # ------------------------------------------------------------------------------------------
    def _step_body(self, x, y, global_batch: int):
        """forward -> fused loss/grad/accuracy -> backward -> (cross-replica reduce +) optimizer."""
        st = self.strategy
        plan = self.fused_plan(x)
        if plan is not None:
            # the whole step -- forward, loss, backward, cross-replica reduction, optimizer -- is ONE kernel
            if self._tracer is not None:
                with self._span("whole step (one cooperative kernel)"):
                    phases = plan.launch_traced(x, y, int(x.shape[0]), global_batch)
                self._tracer.add_kernel_phases("tfde_fused_step", phases)
            else:
                plan.launch(x, y, int(x.shape[0]), 1, global_batch)
            self.optimizer.iterations += 1
            return
        asynchronous = not st.synchronous
        if asynchronous:
            with self._span("parameter-server pull"):
                st.pull(self)              # parameter-server mode: fetch the current weights (N4)
        a = self.arena
        head = self._fusable_head(x)
        if head is not None:
            # classifier head in one launch: Dense + loss/accuracy + d(logits) + dW/db + dh
            self._refresh_w16()
            h = self._prep_input(x)
            for node in self.exec_plan[:-1]:
                with self._span(f"{type(node).__name__}/{node.name}/forward"):
                    h = node.run(h, True)
            dh = head.buf("dx", h.shape) if head.need_dx else None
            with self._span(f"Dense/{head.name}/forward+loss+backward (fused head)"):
                ops.head_fwd_bwd(h, head.kernel.value, head.bias.value if head.bias else None, y,
                                 1.0 / float(global_batch), self.loss.from_probs, False, head.need_dx,
                                 out_dh=dh, out_dw=head.kernel.grad, out_db=head.bias.grad if head.bias else None,
                                 out4=a.grad_tail)
            dy = dh
            for node in reversed(self.exec_plan[:-1]):
                if dy is None:
                    break
                with self._span(f"{type(node).__name__}/{node.name}/backward"):
                    dy = node.run_back(dy)
        else:
            out = self._forward(x, True)
            dgrad = self._in_buf("dloss", out.shape)
            with self._span("loss/softmax cross-entropy + gradient"):
                ops.xent(out, y, 1.0 / float(global_batch), self.loss.from_probs, out_grad=dgrad, out4=a.grad_tail)
            self._backward(dgrad)
        self.join_forks()              # filter-gradient GEMMs forked onto the side stream are done
        if asynchronous:
            with self._span("parameter-server push"):
                st.push(self)          # the owner applies w -= lr*g without waiting for other workers
            a.w[a.n:] = a.g[a.n:]
            a.metrics += a.g[a.n:]
            self.optimizer.iterations += 1
        elif self.strategy.num_replicas_in_sync > 1:
            with self._span("gradient all-reduce + optimizer (fused, peer memory)"):
                self.strategy.comm().reduce_and_apply(a, self.optimizer)
            self.optimizer.iterations += 1
        else:
            with self._span("optimizer"):
                self.optimizer.apply_local(a)

    def fused_plan(self, x=None):
        """The whole-step kernel's plan when this model/strategy/policy is covered by it (Model A topology, fp32,
        synchronous strategy, CUDA) and -- if ``x`` is given -- this batch is (<= 64 rows, uint8/float32)."""
        if not self._fused_checked:
            self._fused_checked = True
            from ..ops import fused_step, mixed
            if (self.device.type == "cuda" and fused_step.enabled() and self.built and self.compiled
                    and mixed.global_policy() == "float32" and self.strategy.synchronous
                    and os.environ.get("TFDE_FUSION", "1") == "1" and not self.loss.from_probs):
                m = fused_step.match(self)
                if m is not None:
                    self._fused = fused_step.FusedStepPlan(self, *m)
        plan = self._fused
        if plan is None or x is None:
            return plan
        if not isinstance(x, torch.Tensor):
            return None
        return plan if plan.supports(x, int(x.shape[0])) else None

    def train_on_batches(self, x, y, steps: Optional[int] = None, global_batch: Optional[int] = None):
        """``steps`` optimizer steps in ONE device launch (Keras ``steps_per_execution``): ``x`` holds the
        per-replica batches of all steps stacked along axis 0 (``[steps*B, ...]`` or ``[steps, B, ...]``), ``y``
        likewise. Each step consumes its own batch and produces its own loss (``last_losses()``). Models the
        whole-step kernel does not cover run the steps one after the other through ``train_on_batch``."""
        if isinstance(x, np.ndarray):
            x = torch.from_numpy(x)
        if isinstance(y, np.ndarray):
            y = torch.from_numpy(y)
        if steps is None:
            if y.dim() != 2:
                raise ValueError("train_on_batches: pass steps=K, or x shaped [K, B, ...] with y shaped [K, B]")
            steps = int(y.shape[0])
            x = x.reshape((-1,) + tuple(x.shape[2:]))
        K = int(steps)
        rows = int(x.shape[0])
        if rows % K:
            raise ValueError(f"{rows} rows do not split into {K} equal steps")
        B = rows // K
        yy = y.reshape(-1)
        if not self.built:
            self.build(tuple(x.shape[1:]))
        if getattr(self.optimizer, "_bound_to", None) is not self:
            self._bind_optimizer()
        gb = global_batch if global_batch is not None else B * self.strategy.num_replicas_in_sync
        plan = self.fused_plan() if self.device.type == "cuda" else None
        if plan is None or not plan.supports(x, B):
            # not covered by the whole-step kernel: the steps run one after the other (each its own launch sequence /
            # graph replay); every step's loss is collected one step late so the reads overlap the next step
            losses = []
            on_host = not x.is_cuda
            for k in range(K):
                self.train_on_batch(x[k * B:(k + 1) * B], yy[k * B:(k + 1) * B], global_batch=gb)
                if k > 0 and on_host:
                    losses.append(self.last_step_loss(lag=1))
            if on_host:
                losses.append(self.last_step_loss())     # device-resident inputs: nothing is read back, no sync
            self._exec_losses_prev, self._exec_losses = getattr(self, "_exec_losses", None), losses
            self._host_tail_mode = "steps"
            return
        self._host_tail_mode = "fused"
        if self._stream is None:
            self._stream = torch.cuda.Stream(device=self.device)
        if x.is_cuda:
            # (conversions, if any, run on the caller's stream BEFORE the edge to the model's stream is recorded; the
            # launch names its stream explicitly: no stream-context switch on the per-launch host path)
            xc = x if x.is_contiguous() else x.contiguous()
            yc = yy if (yy.dtype == torch.int64 and yy.is_contiguous()) else yy.contiguous().long()
            cur = torch.cuda.current_stream(self.device)
            if cur != self._stream:
                self._stream.wait_stream(cur)
            plan.launch(xc, yc, B, K, gb, stream=self._stream)
            self._host_tail = None
        else:
            plan.execute(x, yy, B, K, gb, self._stream)
            self._host_tail = "fused"
        self.optimizer.iterations += K

    def last_losses(self, lag: int = 0):
        """Per-step losses (global mean) of the most recent ``train_on_batches`` / ``train_on_batch`` execution from
        host batches (``lag=1``: of the one before it, so a loop can read every loss while the next execution is
        already running). Waits for that execution only."""
        if getattr(self, "_host_tail_mode", None) == "steps":
            return list((self._exec_losses if lag == 0 else self._exec_losses_prev) or [])
        plan = self._fused
        if plan is not None and getattr(self, "_host_tail", None) == "fused":
            slot = plan.last_slot if lag == 0 else plan.prev_slot
            if slot is None:
                return []
            slot.wait()
            return slot.host_tails[:slot.nsteps, 0].tolist()
        return [self.last_step_loss(lag)]

    def fork_stream(self):
        """Side stream on which layers may run work that only the optimizer consumes (filter gradients)
        concurrently with the rest of backward. None when disabled."""
        if self.device.type != "cuda" or os.environ.get("TFDE_FORK_WGRAD", "1") != "1":
            return None
        if getattr(self, "_side_stream", None) is None:
            self._side_stream = torch.cuda.Stream(device=self.device)
        self._forked = True
        return self._side_stream

    def join_forks(self):
        if getattr(self, "_forked", False):
            torch.cuda.current_stream(self.device).wait_stream(self._side_stream)
            self._forked = False

    def _fusable_head(self, x):
        """The last Dense layer when Dense + loss + its backward can run as the fused head kernel."""
        if os.environ.get("TFDE_FUSION", "1") != "1" or len(self.exec_plan) < 2:
            return None
        last = self.exec_plan[-1]
        if not isinstance(last, L.Dense) or not x.is_cuda:
            return None
        want = "softmax" if self.loss.from_probs else "none"
        if last.activation != want or not ops.head_supported(x, last.kernel.value):
            return None
        return last

    def train_on_batch(self, x, y, global_batch: Optional[int] = None):
        """One optimizer step on this replica's slice of the global batch. Returns nothing (metrics are
        accumulated on device; read them with ``read_metrics()``) so no host sync is forced per step."""
        if not self.compiled:
            raise RuntimeError("compile() the model first")
        if not self.built:
            self.build(tuple(x.shape[1:]))
            self._bind_optimizer()
        elif getattr(self.optimizer, "_bound_to", None) is not self:
            self._bind_optimizer()
        b = int(x.shape[0])
        gb = global_batch if global_batch is not None else b * self.strategy.num_replicas_in_sync
        if self.device.type != "cuda":
            xd = self._to_device(x)
            yd = self._to_device(y).reshape(-1).long()
            self._step_body(xd, yd, gb)
            return
        if self._stream is None:
            self._stream = torch.cuda.Stream(device=self.device)
            ops.native.Workspace.current(self.device)
        if isinstance(x, np.ndarray):
            x = torch.from_numpy(x)
        if isinstance(y, np.ndarray):
            y = torch.from_numpy(y)
        plan = self.fused_plan(x)
        if plan is not None and os.environ.get("TFDE_FUSED_DIRECT", "1") == "1" and self._tracer is None:
            # one kernel per step: no CUDA graph needed. Host batches go through the staging ring (H2D on the copy
            # stream, loss written into mapped pinned memory); device batches are consumed in place.
            yy = y.reshape(-1)
            if x.is_cuda:
                xc = x if x.is_contiguous() else x.contiguous()
                yc = yy if (yy.dtype == torch.int64 and yy.is_contiguous()) else yy.contiguous().long()
                cur = torch.cuda.current_stream(self.device)
                if cur != self._stream:
                    self._stream.wait_stream(cur)
                plan.launch(xc, yc, b, 1, gb, stream=self._stream)
                self._host_tail = None
            else:
                plan.execute(x, yy, b, 1, gb, self._stream)
                self._host_tail = "fused"
            self.optimizer.iterations += 1
            return
        key = (b, tuple(x.shape[1:]), str(x.dtype), gb)
        st = self._graphs.get(key)
        if st is None:
            st = _GraphState()
            xdtype = x.dtype if isinstance(x, torch.Tensor) else torch.from_numpy(x).dtype
            if xdtype not in (torch.uint8, torch.float32):
                xdtype = torch.float32
            st.in_x = torch.empty((b,) + tuple(x.shape[1:]), dtype=xdtype, device=self.device)
            st.in_y = torch.empty((b,), dtype=torch.int64, device=self.device)
            st.slots = [_Slot(tuple(st.in_x.shape), xdtype, b) for _ in range(2)]
            self._graphs[key] = st
        if isinstance(x, np.ndarray):
            x = torch.from_numpy(x)
        if isinstance(y, np.ndarray):
            y = torch.from_numpy(y)
        yy = y.reshape(-1)
        with torch.cuda.stream(self._stream):
            ops.native.Workspace.current(self.device)
            graph_mode = (self.use_cuda_graph and st.eager_steps >= self.graph_warmup_steps
                          and self._tracer is None)
            if graph_mode and not x.is_cuda:
                # Steady state, host inputs: ONE launch per step. The captured graph starts with the H2D
                # copies out of a pinned staging slot and ends with a D2H copy of the 4-float metrics tail
                # into pinned memory; two slots alternate so the host can fill slot k+1 while step k runs.
                slot = st.slots[st.turn & 1]
                st.turn += 1
                if slot.event is not None:
                    slot.event.synchronize()       # the step that last used this slot has finished
                slot.host_x.copy_(x)
                slot.host_y.copy_(yy)
                if slot.graph is None:
                    it0 = self.optimizer.iterations
                    g = torch.cuda.CUDAGraph()
                    zero_copy = os.environ.get("TFDE_ZERO_COPY", "1") == "1"
                    if zero_copy and slot.dev_y is None:
                        try:   # labels are read, and the metrics tail is written, straight through PCIe
                            from ..ops import hostmem
                            slot.dev_y = hostmem.device_alias(slot.host_y, self.device)
                            slot.dev_tail = hostmem.device_alias(slot.host_tail, self.device)
                        except Exception as ex:  # noqa: BLE001  (no UVA alias: fall back to memcpy nodes)
                            log.warning("zero-copy staging unavailable (%s); using memcpy nodes", ex)
                            slot.dev_y = slot.dev_tail = None
                            zero_copy = False
                    slot.overlap = zero_copy and slot.dev_y is not None and os.environ.get("TFDE_OVERLAP_H2D", "1") == "1"
                    if slot.overlap:
                        # the image H2D copy is issued on a copy stream into this slot's own device buffer,
                        # so it overlaps the previous step; the graph waits for it through an event
                        if getattr(self, "_copy_stream", None) is None:
                            self._copy_stream = torch.cuda.Stream(device=self.device)
                        slot.dev_x = torch.empty_like(st.in_x)
                        slot.copy_event = torch.cuda.Event()
                        slot.dev_x.copy_(slot.host_x)
                    with _capture(g, self._stream):
                        if slot.overlap:
                            from ..ops import hostmem
                            self._step_body(slot.dev_x, slot.dev_y, gb)
                            hostmem.copy_small(slot.dev_tail, self.arena.tail)
                        elif zero_copy and slot.dev_y is not None:
                            from ..ops import hostmem
                            st.in_x.copy_(slot.host_x, non_blocking=True)
                            self._step_body(st.in_x, slot.dev_y, gb)
                            hostmem.copy_small(slot.dev_tail, self.arena.tail)
                        else:
                            st.in_x.copy_(slot.host_x, non_blocking=True)
                            st.in_y.copy_(slot.host_y, non_blocking=True)
                            self._step_body(st.in_x, st.in_y, gb)
                            slot.host_tail.copy_(self.arena.tail, non_blocking=True)
                    self.optimizer.iterations = it0   # capture did not execute anything
                    slot.graph = g
                    slot.event = torch.cuda.Event()
                if slot.overlap:
                    with torch.cuda.stream(self._copy_stream):
                        slot.dev_x.copy_(slot.host_x, non_blocking=True)
                        slot.copy_event.record(self._copy_stream)
                    self._stream.wait_event(slot.copy_event)
                slot.graph.replay()
                slot.event.record(self._stream)
                self._host_tail = slot.host_tail
                self._prev_slot, self._last_slot = getattr(self, "_last_slot", None), slot
                self.optimizer.iterations += 1
                return
            st.in_x.copy_(x if x.dtype == st.in_x.dtype else x.to(st.in_x.dtype), non_blocking=True)
            st.in_y.copy_(yy if yy.dtype == torch.int64 else yy.to(torch.int64), non_blocking=True)
            self._host_tail = None
            if graph_mode:
                self.device_step_graph(st, gb).replay()
                self.optimizer.iterations += 1
            else:
                self._step_body(st.in_x, st.in_y, gb)
                st.eager_steps += 1

    def device_step_graph(self, st, gb):
        """CUDA graph of the step alone (inputs already in the static device buffers): used for device
        inputs and by bench.py's device-timed region."""
        if st.graph is None:
            it0 = self.optimizer.iterations
            g = torch.cuda.CUDAGraph()
            with _capture(g, self._stream):
                self._step_body(st.in_x, st.in_y, gb)
            self.optimizer.iterations = it0
            st.graph = g
        return st.graph

    def record_event(self, event):
        """Records ``event`` on the stream this model launches its steps on (device-side timing of a training loop)."""
        if self.device.type == "cuda":
            if self._stream is None:
                self._stream = torch.cuda.Stream(device=self.device)
            event.record(self._stream)
        return event

    def sync(self):
        if self.device.type == "cuda":
            if self._stream is not None:
                self._stream.synchronize()
            else:
                torch.cuda.synchronize(self.device)

    def read_metrics(self, steps: int, reset=True) -> Dict[str, float]:
        """Host read of the device-side running sums accumulated since the last reset (one sync)."""
        self.sync()
        m = self.arena.metrics.detach().cpu().tolist()
        if self.strategy.num_replicas_in_sync > 1:
            self.strategy.comm().check_health()
        if self._fused is not None:
            self._fused.check()
        out = {"loss": m[0] / max(steps, 1)}
        if "accuracy" in self.metrics_names:
            out["accuracy"] = m[1] / max(m[2], 1.0)
        if reset:
            if self._stream is not None:
                with torch.cuda.stream(self._stream):
                    self.arena.metrics.zero_()
            else:
                self.arena.metrics.zero_()
        return out

    def last_step_loss(self, lag: int = 0) -> float:
        """Loss (global mean) of the most recent step, or -- ``lag=1`` -- of the step before it, which lets a
        loop read EVERY step's loss one step late while the next step is already running on the GPU."""
        if getattr(self, "_host_tail", None) == "fused":
            ls = self.last_losses(lag)
            if ls:
                return float(ls[-1])
        elif getattr(self, "_host_tail", None) is not None:
            # already on the host: the graph's last node copied it into the slot's pinned tail
            slot = self._last_slot if lag == 0 else self._prev_slot
            if slot is not None:
                slot.event.synchronize()         # waits for THAT step only; later steps keep running
                return float(slot.host_tail[0])
        self.sync()
        return float(self.arena.tail[0].item())

    # ------------------------------------------------------------------------------------------
    # fit / evaluate / predict
    # ------------------------------------------------------------------------------------------
    def fit(self, x=None, y=None, batch_size=None, epochs=1, steps_per_epoch=None, callbacks=None,
            verbose=1, validation_data=None, initial_epoch=0, shuffle=True):
        from ..data.dataset import Dataset, DistributedDataset, as_distributed
        if not self.compiled:
            raise RuntimeError("compile() the model first")
        if not isinstance(x, (Dataset, DistributedDataset)):
            if y is None:
                raise ValueError("fit(x, y): y is required when x is an array")
            n_rep = self.strategy.num_replicas_in_sync
            gbs = (batch_size or 32) * 1
            ds = Dataset.from_tensor_slices((np.asarray(x), np.asarray(y)))
            if shuffle:
                ds = ds.shuffle(min(len(x), 10000), seed=self.seed)
            x = ds.batch(gbs)
            del n_rep
        dist_ds = as_distributed(x, self.strategy)
        history = History()
        cbs = CallbackList([history] + list(callbacks or []), self)
        self.stop_training = False
        cbs.on_train_begin()
        it = iter(dist_ds)
        chief_verbose = verbose and self.strategy.is_chief
        for epoch in range(initial_epoch, epochs):
            cbs.on_epoch_begin(epoch)
            t0 = time.time()
            steps = 0
            while steps_per_epoch is None or steps < steps_per_epoch:
                try:
                    xb, yb, gb = next(it)
                except StopIteration:
                    if steps_per_epoch is None:
                        it = iter(dist_ds)   # next epoch restarts the dataset
                        break
                    if steps == 0 and epoch > initial_epoch:
                        it = iter(dist_ds)
                        continue
                    log.warning("input ran out of data after %d steps of epoch %d", steps, epoch + 1)
                    self.stop_training = True
                    break
                group = [(xb, yb)]
                want = self.steps_per_execution
                if steps_per_epoch is not None:
                    want = min(want, steps_per_epoch - steps)
                pending = None
                while len(group) < want:
                    # steps_per_execution: hand several batches to the device at once (same shape only)
                    try:
                        xn, yn, gn = next(it)
                    except StopIteration:
                        break
                    if gn != gb or tuple(xn.shape) != tuple(xb.shape):
                        pending = (xn, yn, gn)
                        break
                    group.append((xn, yn))
                cbs.on_train_batch_begin(steps)
                if len(group) == 1:
                    self.train_on_batch(xb, yb, global_batch=gb)
                else:
                    as_t = lambda a: a if isinstance(a, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(a))
                    xs = torch.cat([as_t(g[0]) for g in group], 0)
                    ys = torch.cat([as_t(g[1]).reshape(-1) for g in group], 0)
                    self.train_on_batches(xs, ys, steps=len(group), global_batch=gb)
                steps += len(group)
                cbs.on_train_batch_end(steps - 1)
                if pending is not None:
                    # a batch of another shape ended the group: it is its own step
                    self.train_on_batch(pending[0], pending[1], global_batch=pending[2])
                    steps += 1
            logs = self.read_metrics(steps)
            dt = time.time() - t0
            if validation_data is not None:
                val = self.evaluate(*validation_data, batch_size=batch_size, verbose=0, return_dict=True)
                logs.update({f"val_{k}": v for k, v in val.items()})
            if chief_verbose:
                extra = " - ".join(f"{k}: {v:.4f}" for k, v in logs.items())
                print(f"Epoch {epoch + 1}/{epochs}\n{steps}/{steps} - {dt:.2f}s - "
                      f"{1e3 * dt / max(steps, 1):.2f}ms/step - {extra}", flush=True)
            cbs.on_epoch_end(epoch, logs)
            if self.stop_training:
                break
        cbs.on_train_end()
        return history

    def _eval_replicas(self):
        """[self] + one replica of this model per additional local device of a LOCAL MirroredStrategy (in-process
        mirrored evaluation, mnist_keras_distributed.py:243). Replicas are built once and refreshed from this
        model's weights at every evaluate()."""
        devs = getattr(self.strategy, "local_devices", None)
        if not devs or len(devs) < 2 or self.device.type != "cuda" or os.environ.get("TFDE_MIRRORED_EVAL", "1") != "1":
            return [self]
        reps = getattr(self, "_replicas", None)
        if reps is None or [r.device for r in reps] != list(devs[1:]):
            from ..parallel.strategy import Strategy
            reps = []
            cur = torch.cuda.current_device()
            try:
                for d in devs[1:]:
                    with Strategy(device=str(d)).scope():
                        layers = [L.layer_from_config(type(l).__name__, l.get_config()) for l in self.layers]
                        rep = Sequential(layers, name=f"{self.name}_replica_{d.index}", seed=self.seed)
                    rep.build(self.input_shape)
                    rep.loss, rep.metrics_names, rep.compiled = self.loss, self.metrics_names, False
                    reps.append(rep)
            finally:
                torch.cuda.set_device(cur)
            self._replicas = reps
        self.sync()
        for rep in reps:
            rep.arena.w.copy_(self.arena.w)
            rep.state.copy_(self.state)
        torch.cuda.synchronize(self.device)
        return [self] + reps

    def evaluate(self, x=None, y=None, batch_size=None, steps=None, verbose=0, return_dict=False):
        """Distributed evaluation (the MirroredStrategy eval of mnist_keras_distributed.py:243): each
        replica scores its slice of every batch; (loss_sum, correct, count) are SUM-reduced (G14). Replicas are
        the processes of the strategy's group and/or -- local MirroredStrategy -- the GPUs of this process."""
        from ..data.dataset import Dataset, DistributedDataset, as_distributed
        if not isinstance(x, (Dataset, DistributedDataset)):
            ds = Dataset.from_tensor_slices((np.asarray(x), np.asarray(y))).batch(batch_size or 32)
            x = ds
        dist_ds = as_distributed(x, self.strategy, for_eval=True)
        self.sync()
        replicas = None
        acc = None
        n_steps = 0
        for xb, yb, gb in dist_ds:
            if steps is not None and n_steps >= steps:
                break
            n = int(xb.shape[0])
            if n > 0:
                if not self.built:
                    self.build(tuple(xb.shape[1:]))
                if replicas is None:
                    replicas = self._eval_replicas()
                    acc = [torch.zeros(4, dtype=torch.float32, device=r.device) for r in replicas]
                R = len(replicas)
                for i, rep in enumerate(replicas):
                    lo, hi = n * i // R, n * (i + 1) // R
                    if hi <= lo:
                        continue
                    ctx = torch.cuda.device(rep.device) if rep.device.type == "cuda" else contextlib.nullcontext()
                    with ctx:
                        xd = rep._to_device(xb[lo:hi])
                        yd = rep._to_device(yb[lo:hi]).reshape(-1).long()
                        out = rep._forward(xd, False)
                        o4, _ = ops.xent(out, yd, 1.0, self.loss.from_probs if self.loss else False,
                                         out_grad=rep._in_buf("dloss_eval", out.shape))
                        acc[i] += o4.detach()          # stays on the device: no host sync per batch
            n_steps += 1
        tot = torch.zeros(4, dtype=torch.float64)
        for a in acc or []:
            tot += a.to("cpu", torch.float64)
        red = tot.to(torch.float32).to(self.device)
        self.strategy.all_reduce_sum_(red)
        r = red.cpu().tolist()
        count = max(r[2], 1.0)
        res = {"loss": r[0] / count}
        if "accuracy" in self.metrics_names or not self.compiled:
            res["accuracy"] = r[1] / count
        if verbose and self.strategy.is_chief:
            print(" - ".join(f"{k}: {v:.4f}" for k, v in res.items()), flush=True)
        return res if return_dict else [res[k] for k in res]

    def predict(self, x, batch_size=32):
        xs = np.asarray(x) if not isinstance(x, torch.Tensor) else x
        outs = []
        for i in range(0, len(xs), batch_size):
            xb = self._to_device(xs[i:i + batch_size])
            if not self.built:
                self.build(tuple(xb.shape[1:]))
            outs.append(self._forward(xb, False).detach().cpu().clone())
        return torch.cat(outs, 0).numpy()

    # ------------------------------------------------------------------------------------------
    # introspection / weights
    # ------------------------------------------------------------------------------------------
    def count_params(self):
        return sum(v.size for v in self.variables)

    def summary(self, print_fn=print):
        if not self.built:
            self.build()
        lines = [f'Model: "{self.name}"', "_" * 65,
                 f"{'Layer (type)':<29}{'Output Shape':<26}{'Param #':<10}", "=" * 65]
        for l in self.layers:
            shape = "(None, " + ", ".join(str(s) for s in l.out_shape) + ")"
            lines.append(f"{(l.name + ' (' + type(l).__name__ + ')'):<29}{shape:<26}{l.count_params():<10}")
            lines.append("_" * 65)
        train = sum(v.size for v in self.variables if v.trainable)
        total = self.count_params()
        lines[-1] = "=" * 65
        lines += [f"Total params: {total:,}", f"Trainable params: {train:,}",
                  f"Non-trainable params: {total - train:,}", "_" * 65]
        text = "\n".join(lines)
        print_fn(text)
        return text

    @property
    def trainable_variables(self):
        return [v for v in self.variables if v.trainable]

    @property
    def non_trainable_variables(self):
        return [v for v in self.variables if not v.trainable]

    def get_weights(self) -> List[np.ndarray]:
        """Keras order: per layer, in creation order (kernel, bias | gamma, beta, moving_mean, moving_variance)."""
        self.sync()
        return [v.numpy() for v in self.variables]

    def set_weights(self, weights: Sequence[np.ndarray]):
        if len(weights) != len(self.variables):
            raise ValueError(f"expected {len(self.variables)} arrays, got {len(weights)}")
        for v, w in zip(self.variables, weights):
            w = np.asarray(w, dtype=np.float32)
            if tuple(w.shape) != v.shape:
                raise ValueError(f"{v.name}: expected shape {v.shape}, got {w.shape}")
            v.value.copy_(torch.from_numpy(w).to(self.device))

    def named_weights(self) -> Dict[str, np.ndarray]:
        self.sync()
        return {v.name: v.numpy() for v in self.variables}

    def save_weights(self, path: str):
        from ..ckpt.checkpoint import save_weights
        save_weights(self, path)

    def load_weights(self, path: str):
        from ..ckpt.checkpoint import load_weights
        load_weights(self, path)

    @property
    def graph_captured(self) -> bool:
        """True once a CUDA graph of the training step exists (either flavour)."""
        for v in self._graphs.values():
            if isinstance(v, _GraphState) and (v.graph is not None or any(sl.graph is not None for sl in v.slots or [])):
                return True
        return False

    def invalidate_graphs(self):
        """Drop captured CUDA graphs (needed only if buffers were re-allocated; weight updates through
        set_weights/load_weights keep all addresses and need no re-capture)."""
        for v in self._graphs.values():
            if isinstance(v, _GraphState):
                v.graph = None
                v.eager_steps = 0
                for sl in v.slots or []:
                    sl.graph = None
