"""Binding of the whole-step kernel (``csrc/kernels/fused_step.cu``): forward, loss, backward, cross-replica
gradient reduction and optimizer of the reference's Keras MNIST CNN (distributed_with_keras.py:33-43) in ONE
cooperative launch -- optionally ``steps_per_execution`` steps per launch.

:class:`FusedStepPlan` owns everything the kernel needs besides the model's arenas: the split-K / conv partials,
the barrier state, and -- under a multi-replica strategy -- the symmetric (peer-mapped) receive buffers and
flags of the in-kernel all-reduce. It also owns the end-to-end staging ring (device input slots, mapped pinned
loss tails, copy stream, events) used by ``Sequential.train_on_batches``.
"""
from __future__ import annotations

import ctypes
import logging
import os
from typing import Optional

import torch

from . import native

log = logging.getLogger("tfde_b200")

MAX_RANKS = 8
MAX_B = 64
MAX_CLASSES = 16


class _FusedStepDesc(ctypes.Structure):
    _fields_ = [
        ("x", ctypes.c_void_p), ("y", ctypes.c_void_p),
        ("x_is_u8", ctypes.c_int), ("in_scale", ctypes.c_float), ("x_step_bytes", ctypes.c_long),
        ("B", ctypes.c_int), ("nsteps", ctypes.c_int), ("ncls", ctypes.c_int), ("opt", ctypes.c_int),
        ("w", ctypes.c_void_p), ("g", ctypes.c_void_p), ("s1", ctypes.c_void_p), ("s2", ctypes.c_void_p),
        ("off_ck", ctypes.c_long), ("off_cb", ctypes.c_long), ("off_dk", ctypes.c_long), ("off_db", ctypes.c_long),
        ("off_hk", ctypes.c_long), ("off_hb", ctypes.c_long), ("n", ctypes.c_long),
        ("loss_scale", ctypes.c_float), ("rank", ctypes.c_int), ("world", ctypes.c_int), ("use_tc", ctypes.c_int),
        ("hp", ctypes.c_void_p), ("metrics", ctypes.c_void_p), ("tails", ctypes.c_void_p),
        ("part", ctypes.c_void_p), ("hbuf", ctypes.c_void_p), ("cpart", ctypes.c_void_p),
        ("small_block", ctypes.c_void_p),
        ("state", ctypes.c_void_p), ("status", ctypes.c_void_p),
        ("timeout_ns", ctypes.c_ulonglong),
        ("rs", ctypes.c_void_p * MAX_RANKS), ("gsum", ctypes.c_void_p * MAX_RANKS),
        ("srecv", ctypes.c_void_p * MAX_RANKS), ("flags", ctypes.c_void_p * MAX_RANKS),
        ("prof", ctypes.c_void_p),
        ("ready", ctypes.c_void_p), ("ready_tag", ctypes.c_uint),
    ]


_sz = ctypes.c_size_t
for _n, _a, _r in [
    ("tfde_fused_step_desc_bytes", [], _sz), ("tfde_fused_step_part_floats", [], _sz),
    ("tfde_fused_step_hbuf_floats", [], _sz), ("tfde_fused_step_cpart_floats", [], _sz),
    ("tfde_fused_step_small_floats", [], _sz), ("tfde_fused_step_state_words", [], _sz),
    ("tfde_fused_step_rs_floats", [ctypes.c_int], _sz), ("tfde_fused_step_gsum_floats", [], _sz),
    ("tfde_fused_step_srecv_floats", [ctypes.c_int], _sz), ("tfde_fused_step_flag_words", [ctypes.c_int], _sz),
    ("tfde_fused_step_blocks", [], ctypes.c_int), ("tfde_fused_step_prof_stamps", [], ctypes.c_int),
    ("tfde_fused_step", [ctypes.POINTER(_FusedStepDesc), ctypes.c_void_p], ctypes.c_int),
    ("tfde_fused_exec", [ctypes.POINTER(_FusedStepDesc), ctypes.c_void_p, ctypes.c_void_p, _sz, _sz,
                         ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p], ctypes.c_int),
    ("tfde_fused_exec_streamed", [ctypes.POINTER(_FusedStepDesc), ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                  _sz, _sz, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p],
     ctypes.c_int),
    ("tfde_event_create", [ctypes.POINTER(ctypes.c_void_p)], ctypes.c_int),
    ("tfde_event_destroy", [ctypes.c_void_p], ctypes.c_int),
    ("tfde_event_synchronize", [ctypes.c_void_p], ctypes.c_int),
    ("tfde_event_query", [ctypes.c_void_p], ctypes.c_int),
]:
    native.register_optional(_n, _a, _r)


def enabled() -> bool:
    return os.environ.get("TFDE_FUSED_STEP", "1") == "1"


def streamed_inputs() -> bool:
    """End-to-end executions launch the kernel without waiting for the host->device copies; the kernel takes up each
    step when that step's ready tag has landed (``TFDE_FUSED_STREAM=0``: all copies first, then the launch)."""
    return os.environ.get("TFDE_FUSED_STREAM", "1") == "1"


def match(model):
    """(conv_node, dense, head) when ``model`` is the topology the whole-step kernel covers, else None:
    Conv2D(32, 3, relu, valid, Cin=1, 28x28) + MaxPooling2D(2) [already fused by the model's fusion pass]
    -> Flatten -> Dense(64, relu) -> Dense(n <= 16), all with biases."""
    from ..keras import layers as L
    plan = model.exec_plan
    if len(plan) != 4:
        return None
    conv, flat, dense, head = plan
    if not (isinstance(conv, L.FusedConvReluPool) and isinstance(flat, L.Flatten)
            and isinstance(dense, L.Dense) and isinstance(head, L.Dense)):
        return None
    if tuple(conv.in_shape) != (28, 28, 1) or tuple(conv.out_shape) != (13, 13, 32):
        return None
    if dense.units != 64 or dense.activation != "relu" or dense.bias is None:
        return None
    if head.units > MAX_CLASSES or head.activation != "none" or head.bias is None:
        return None
    if tuple(dense.kernel.shape) != (5408, 64):
        return None
    return conv, dense, head


class _Slot:
    """One in-flight execution: device inputs for up to K steps, mapped pinned loss tails, events."""

    def __init__(self, plan: "FusedStepPlan", K: int, B: int, x_dtype):
        from . import hostmem
        L = native.load()
        dev = plan.device
        self.K, self.B = K, B
        self.x = torch.empty((K, B, 28, 28, 1), dtype=x_dtype, device=dev)
        self.y = torch.empty((K, B), dtype=torch.int64, device=dev)
        self.host_tails = torch.zeros((K, 4), dtype=torch.float32).pin_memory()
        self.dev_tails = hostmem.device_alias(self.host_tails, dev)
        # pinned staging for callers whose batches are not pinned
        self.host_x = None
        self.host_y = None
        # streamed inputs: per-step ready words on the device, their pinned source, a tag that grows with every use
        self.ready = torch.zeros(K, dtype=torch.int32, device=dev)
        self.host_tags = torch.zeros(K, dtype=torch.int32).pin_memory()
        self.tag = 0
        ev = ctypes.c_void_p()
        native.check(L.tfde_event_create(ctypes.byref(ev)), "event_create", kernels=0)
        self.ev_copied = ev.value
        ev = ctypes.c_void_p()
        native.check(L.tfde_event_create(ctypes.byref(ev)), "event_create", kernels=0)
        self.ev_done = ev.value
        self.busy = False
        self.nsteps = 0
        # the zero fill of `ready` ran on the caller's current stream; the copy stream writes tags into it and the
        # kernel (another stream) polls it: it must have landed before either starts
        torch.cuda.current_stream(dev).synchronize()

    def wait(self):
        if self.busy:
            native.check(native.load().tfde_event_synchronize(self.ev_done), "event_synchronize", kernels=0)
            self.busy = False


class FusedStepPlan:
    def __init__(self, model, conv, dense, head):
        L = native.load()
        if ctypes.sizeof(_FusedStepDesc) != L.tfde_fused_step_desc_bytes():
            raise RuntimeError("fused-step descriptor layout mismatch between Python and the native library")
        self.model = model
        self.device = dev = model.device
        st = model.strategy
        self.world, self.rank = st.num_replicas_in_sync, st.replica_id
        if self.world > MAX_RANKS:
            raise ValueError(f"whole-step kernel supports up to {MAX_RANKS} replicas")
        a = model.arena
        f32 = dict(dtype=torch.float32, device=dev)
        self.part = torch.empty(L.tfde_fused_step_part_floats(), **f32)
        self.hbuf = torch.empty(L.tfde_fused_step_hbuf_floats(), **f32)
        self.cpart = torch.empty(L.tfde_fused_step_cpart_floats(), **f32)
        self.small = torch.zeros(L.tfde_fused_step_small_floats(), **f32)
        self.state = torch.zeros(L.tfde_fused_step_state_words(), dtype=torch.int32, device=dev)
        self.status = torch.zeros(4, dtype=torch.int32, device=dev)
        d = self.desc = _FusedStepDesc()
        d.w, d.g = a.w.data_ptr(), a.g.data_ptr()
        ck, cb = conv.conv.kernel, conv.conv.bias
        d.off_ck, d.off_cb = ck.offset, cb.offset
        d.off_dk, d.off_db = dense.kernel.offset, dense.bias.offset
        d.off_hk, d.off_hb = head.kernel.offset, head.bias.offset
        d.n, d.ncls = a.n, head.units
        d.part, d.hbuf, d.cpart = self.part.data_ptr(), self.hbuf.data_ptr(), self.cpart.data_ptr()
        d.small_block = self.small.data_ptr()
        d.state, d.status = self.state.data_ptr(), self.status.data_ptr()
        d.metrics = a.metrics.data_ptr()
        d.rank, d.world = self.rank, self.world
        # GEMM precision of the three per-CTA products: exact fp32 FFMA (default; what the reference's fp32 graph asks
        # for) or -- experimental, TFDE_FUSED_TC=1 + TFDE_FUSED_TC_FORCE=1 -- TF32 tcgen05.mma with TMEM accumulators.
        # Measured: 27.1 vs 28.4 us per step (the step is bound by its barriers, not by the GEMMs), and the CUDA
        # runtime grants kernels that allocate tensor memory ONE CTA per SM, so a cooperative launch of the 169-CTA
        # grid is refused; the forced variant launches plainly and relies on the bounded waits. Hence opt-in.
        self.tensor_core = os.environ.get("TFDE_FUSED_TC", "0") == "1"
        d.use_tc = 1 if self.tensor_core else 0
        d.timeout_ns = int(float(os.environ.get("TFDE_P2P_TIMEOUT_S", "20")) * 1e9)
        self.in_scale_u8 = float(conv.in_scale)
        self._buffers = []
        if self.world > 1:
            comm = st.comm()
            W = self.world
            self.rs = comm.symm_buffer(4 * L.tfde_fused_step_rs_floats(W), "fs_rs")
            self.gsum = comm.symm_buffer(4 * L.tfde_fused_step_gsum_floats(), "fs_gsum")
            self.srecv = comm.symm_buffer(4 * L.tfde_fused_step_srecv_floats(W), "fs_srecv")
            self.flags = comm.symm_buffer(4 * L.tfde_fused_step_flag_words(W), "fs_flags")
            for r in range(W):
                d.rs[r], d.gsum[r] = self.rs.peer_ptrs[r], self.gsum.peer_ptrs[r]
                d.srecv[r], d.flags[r] = self.srecv.peer_ptrs[r], self.flags.peer_ptrs[r]
            import torch.distributed as dist
            torch.cuda.synchronize(dev)
            dist.barrier(group=st.group)          # everyone has mapped everything before anyone launches
            log.info("whole-step kernel: in-kernel all-reduce over %d replicas (peer memory)", W)
        # the zero fills above ran on the caller's current stream, the kernel runs on the model's: one-time join
        torch.cuda.synchronize(dev)
        self._slots = {}
        self._pin_cache = {}
        self._turn = 0
        self._copy_stream = None
        self.last_slot: Optional[_Slot] = None
        self.prev_slot: Optional[_Slot] = None

    # ------------------------------------------------------------------------------------------
    def _fill(self, x, y, B, nsteps, global_batch, tails):
        m, d = self.model, self.desc
        opt = m.optimizer
        opt.ensure_slots(m.arena)                  # replicated optimizer state (full arena) on every rank
        m.arena.state_sharded = False
        m.arena.state_slice = (0, m.arena.n_total)
        d.x, d.y = x.data_ptr(), y.data_ptr()
        d.x_is_u8 = 1 if x.dtype == torch.uint8 else 0
        d.in_scale = self.in_scale_u8 if x.dtype == torch.uint8 else 1.0
        d.x_step_bytes = B * 784 * x.element_size()
        d.B, d.nsteps, d.opt = B, nsteps, opt.native_code
        d.s1 = m.arena.s1.data_ptr() if m.arena.s1 is not None else None
        d.s2 = m.arena.s2.data_ptr() if m.arena.s2 is not None else None
        d.loss_scale = 1.0 / float(global_batch)
        d.hp = opt.hparams.data_ptr()
        d.tails = tails.data_ptr() if tails is not None else None

    def supports(self, x, B: int) -> bool:
        return (1 <= B <= MAX_B and x.dtype in (torch.uint8, torch.float32)
                and not self.model.loss.from_probs and self.model.strategy.synchronous)

    def _pinned(self, t: torch.Tensor) -> bool:
        """``t.is_pinned()`` costs a driver query (~10 us) per call: remembered per host allocation. A stale entry (the
        address re-used by pageable memory) only costs asynchrony: cudaMemcpyAsync stages pageable sources itself."""
        key = (t.data_ptr(), t.numel() * t.element_size())
        hit = self._pin_cache.get(key)
        if hit is None:
            if len(self._pin_cache) > 256:
                self._pin_cache.clear()
            hit = self._pin_cache[key] = bool(t.is_pinned())
        return hit

    def launch(self, x: torch.Tensor, y: torch.Tensor, B: int, nsteps: int, global_batch: int, tails=None, stream=None):
        """``nsteps`` optimizer steps on device-resident inputs x [nsteps*B, 28, 28, 1] / y [nsteps*B] (int64),
        on ``stream`` (default: the current stream)."""
        self._fill(x, y, B, nsteps, global_batch, tails)
        sp = stream.cuda_stream if stream is not None else native.stream_ptr()
        native.check(native.load().tfde_fused_step(ctypes.byref(self.desc), sp), "fused_step")

    # ------------------------------------------------------------------------------------------
    def execute(self, x, y, B: int, nsteps: int, global_batch: int, stream: torch.cuda.Stream) -> _Slot:
        """End-to-end execution from HOST batches: per-step H2D copies from pinned memory on the copy stream, one
        launch for ``nsteps`` steps, every step's [loss, correct, rows] written into mapped pinned memory."""
        L = native.load()
        xdtype = x.dtype
        key = (nsteps, B, xdtype)
        slots = self._slots.get(key)
        if slots is None:
            slots = self._slots[key] = [_Slot(self, nsteps, B, xdtype) for _ in range(2)]
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(device=self.device)
        slot = slots[self._turn & 1]
        self._turn += 1
        slot.wait()                                   # the execution that last used this slot has finished
        if not (self._pinned(x) and x.is_contiguous()):
            if slot.host_x is None:
                slot.host_x = torch.empty((nsteps * B, 28, 28, 1), dtype=xdtype).pin_memory()
            slot.host_x.view(-1).copy_(x.reshape(-1))
            x = slot.host_x
        if y.dtype != torch.int64 or not (self._pinned(y) and y.is_contiguous()):
            if slot.host_y is None:
                slot.host_y = torch.empty((nsteps * B,), dtype=torch.int64).pin_memory()
            slot.host_y.copy_(y.reshape(-1))
            y = slot.host_y
        self._fill(slot.x, slot.y, B, nsteps, global_batch, slot.dev_tails)
        if streamed_inputs() and hasattr(L, "tfde_fused_exec_streamed"):
            slot.tag = (slot.tag + 1) & 0xFFFFFFFF     # (slot.wait() above: nobody reads the previous tags any more)
            slot.host_tags.fill_(slot.tag - (1 << 32) if slot.tag >= (1 << 31) else slot.tag)   # uint32 bit pattern
            self.desc.ready, self.desc.ready_tag = slot.ready.data_ptr(), slot.tag
            try:
                native.check(L.tfde_fused_exec_streamed(ctypes.byref(self.desc), x.data_ptr(), y.data_ptr(),
                                                        slot.host_tags.data_ptr(), B * 784 * x.element_size(), B * 8,
                                                        self._copy_stream.cuda_stream, slot.ev_copied, slot.ev_done,
                                                        stream.cuda_stream), "fused_exec")
            finally:
                self.desc.ready, self.desc.ready_tag = None, 0
        else:
            native.check(L.tfde_fused_exec(ctypes.byref(self.desc), x.data_ptr(), y.data_ptr(),
                                           B * 784 * x.element_size(), B * 8, self._copy_stream.cuda_stream,
                                           slot.ev_copied, slot.ev_done, stream.cuda_stream), "fused_exec")
        slot.busy = True
        slot.nsteps = nsteps
        slot.src = (x, y)                              # keep the pinned sources alive until the copies ran
        self.prev_slot, self.last_slot = self.last_slot, slot
        return slot

    PHASES = ["P1 conv+pool+split-K partial", "B1 grid barrier", "P2 row reduce + head (64 CTAs)", "B2 grid barrier",
              "P3 load dh (+h, dz)", "P3 head grads (16 CTAs)", "P3 dW/dx/conv partial", "fence + two-level ticket",
              "tail (winner CTA)", "slice exchange/update", "wait step done"]

    def profile(self, x, y, B, nsteps, global_batch):
        """Per-phase device time of the LAST step of one launch: globaltimer stamps taken by thread 0 of every CTA.
        Returns {phase: (mean_us, max_us over CTAs)} plus the step's span. Debug/profiling aid (tools/fused_step_phases.py)."""
        L = native.load()
        ns = L.tfde_fused_step_prof_stamps()
        nb = L.tfde_fused_step_blocks()
        prof = torch.zeros((nb, ns), dtype=torch.int64, device=self.device)
        self.desc.prof = prof.data_ptr()
        try:
            self.launch(x, y, B, nsteps, global_batch)
            torch.cuda.synchronize(self.device)
        finally:
            self.desc.prof = None
        t = prof.cpu().double()
        out = {}
        for i, name in enumerate(self.PHASES):
            d = (t[:, i + 1] - t[:, i]) / 1e3
            out[name] = (float(d.mean()), float(d.max()))
        win = int(torch.argmax(t[:, 12]))            # the ticket winner is the only CTA that stamps inside the tail
        if t[win, 12] > 0:
            out["tail: local reduce of conv partials"] = (float((t[win, 12] - t[win, 8]) / 1e3),) * 2
            if t[win, 13] > 0:
                out["tail: pack + push small block to peers"] = (float((t[win, 13] - t[win, 12]) / 1e3),) * 2
                out["tail: gather peers' small blocks (wait)"] = (float((t[win, 14] - t[win, 13]) / 1e3),) * 2
                out["tail: sum + optimizer + release"] = (float((t[win, 9] - t[win, 14]) / 1e3),) * 2
            else:
                out["tail: optimizer + release"] = (float((t[win, 9] - t[win, 12]) / 1e3),) * 2
        out["step span (first start -> last end)"] = (float((t[:, 10].max() - t[:, 0].min()) / 1e3),) * 2
        out["P1..P3 critical path (max end - min start)"] = (float((t[:, 7].max() - t[:, 0].min()) / 1e3),) * 2
        return out

    def launch_traced(self, x, y, B, global_batch):
        """One step with the in-kernel phase stamps on: returns [(phase, start_us, dur_us)] for CTA-averaged phases
        relative to the kernel's first stamp (ProfilerHook timelines)."""
        L = native.load()
        prof = torch.zeros((L.tfde_fused_step_blocks(), L.tfde_fused_step_prof_stamps()), dtype=torch.int64,
                           device=self.device)
        self.desc.prof = prof.data_ptr()
        try:
            self.launch(x, y, B, 1, global_batch)
            torch.cuda.current_stream(self.device).synchronize()
        finally:
            self.desc.prof = None
        t = prof.cpu().double()
        t0 = float(t[:, 0].min())
        out = []
        for i, name in enumerate(self.PHASES):
            start = float(t[:, i].mean() - t0) / 1e3
            out.append((name, start, max(float((t[:, i + 1] - t[:, i]).mean()) / 1e3, 0.0)))
        return out

    def check(self):
        st = self.status.cpu().tolist()
        if st[0] != 0:
            what = {10: "grid barrier", 11: "step-complete flag", 12: "streamed input batch (host->device copy)", 21: "small-tensor exchange with peer",
                    22: "reduce-scatter contribution from peer", 23: "reduced slice from owner"}.get(st[0], "wait")
            raise RuntimeError(
                f"whole-step kernel timed out on rank {self.rank}: {what} (code {st[0]}), epoch/target {st[1]}, "
                f"peer/CTA {st[2]}, slice/phase {st[3]} (peer dead or diverged; restart from the last checkpoint)")
