"""Gradient communication backends.

* :class:`P2PComm` (CUDA): symmetric heap + the fused all-reduce/optimizer kernel
  (``csrc/kernels/allreduce_p2p.cu``). ``torch.distributed`` is used once, at set-up, to exchange CUDA
  IPC handles; no NCCL collective runs per step.
* :class:`GlooComm` (CPU): ``dist.all_reduce`` + the reference optimizer -- the plumbing/test path of
  BASELINE.json config 1 ("2 workers via TF_CONFIG on localhost CPU/gloo").

Both implement ``reduce_and_apply(arena)`` = SUM-reduce the flat gradient arena across replicas
(including its 4-float metrics tail, G14) and apply the optimizer to the flat parameter arena, which
is the whole of what TF's ``CollectiveReduce`` + ``apply_gradients`` do per step
(distributed_with_keras.py:63).
"""
from __future__ import annotations

import ctypes
import logging
import os
from typing import List, Optional

import torch
import torch.distributed as dist

from ..ops import native
from ..ops import reference as ref

log = logging.getLogger("tfde_b200")

MAX_RANKS = 8
MAX_BLOCKS = 256
ONE_SHOT, TWO_SHOT, NVLS = 0, 1, 2
_MODE_NAMES = {ONE_SHOT: "one_shot", TWO_SHOT: "two_shot", NVLS: "nvls"}


def make_comm(strategy):
    if strategy.device.type == "cuda":
        return P2PComm(strategy)
    return GlooComm(strategy)


# ----------------------------------------------------------------------------------------------
# flat arena shared by model / optimizer / comm
# ----------------------------------------------------------------------------------------------
class Arena:
    """The flat buffers one model trains with. ``n`` trainable floats + ``n_extra`` metric scalars."""

    N_EXTRA = 4

    def __init__(self, n: int, device, comm=None):
        self.n = int(n)
        self.n_total = ((self.n + self.N_EXTRA + 3) // 4) * 4
        self.n_extra = self.n_total - self.n
        self.device = torch.device(device)
        self.comm = comm
        if comm is not None and isinstance(comm, P2PComm) and comm.world > 1:
            # with NVLS both arenas are multicast-bound: gradients are read, parameters written, through
            # the switch (allreduce_nvls_kernel); without it only w needs peer mappings (two-shot all-gather)
            self.w = comm.symm_empty(self.n_total, "w", multicast=True)
            self.g = (comm.symm_empty(self.n_total, "g", multicast=True) if comm.nvls
                      else torch.zeros(self.n_total, dtype=torch.float32, device=self.device))
        else:
            self.w = torch.zeros(self.n_total, dtype=torch.float32, device=self.device)
            self.g = torch.zeros(self.n_total, dtype=torch.float32, device=self.device)
        self.metrics = torch.zeros(self.n_extra, dtype=torch.float32, device=self.device)
        self.s1: Optional[torch.Tensor] = None
        self.s2: Optional[torch.Tensor] = None
        self.w_bf16: Optional[torch.Tensor] = None
        self.state_sharded = False           # True when optimizer slots hold only this rank's slice
        self.state_slice = (0, self.n_total)  # (offset, length) of the slots within the arena

    @property
    def tail(self) -> torch.Tensor:
        """[loss*scale, correct, rows, 0] of the last step after reduction (lives at w[n:])."""
        return self.w[self.n:self.n + 4]

    @property
    def grad_tail(self) -> torch.Tensor:
        return self.g[self.n:self.n + 4]


class CommBase:
    def __init__(self, strategy):
        self.strategy = strategy
        self.rank, self.world = strategy.rank, strategy.world_size
        self.device = strategy.device

    def broadcast_(self, t, src=0):
        raise NotImplementedError

    def all_reduce_sum_(self, t):
        raise NotImplementedError

    def prepare(self, arena: Arena, optimizer):
        """Called once after the arenas exist and the optimizer is known."""

    def reduce_and_apply(self, arena: Arena, optimizer):
        raise NotImplementedError

    def check_health(self):
        pass

    def close(self):
        pass


# ----------------------------------------------------------------------------------------------
# CPU / gloo
# ----------------------------------------------------------------------------------------------
class GlooComm(CommBase):
    def broadcast_(self, t, src=0):
        dist.broadcast(t, src=src)
        return t

    def all_reduce_sum_(self, t):
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return t

    def reduce_and_apply(self, arena: Arena, optimizer):
        if self.world > 1:
            dist.all_reduce(arena.g, op=dist.ReduceOp.SUM)
        optimizer.apply_local(arena)


# ----------------------------------------------------------------------------------------------
# CUDA / peer-to-peer
# ----------------------------------------------------------------------------------------------
class _ARDesc(ctypes.Structure):
    _fields_ = [
        ("rank", ctypes.c_int), ("world", ctypes.c_int), ("mode", ctypes.c_int), ("opt", ctypes.c_int),
        ("wire_bf16", ctypes.c_int), ("n_extra", ctypes.c_int), ("blocks", ctypes.c_int),
        ("reserved", ctypes.c_int),
        ("n", ctypes.c_long), ("n_total", ctypes.c_long), ("slot_floats", ctypes.c_long),
        ("g", ctypes.c_void_p), ("w", ctypes.c_void_p), ("s1", ctypes.c_void_p), ("s2", ctypes.c_void_p),
        ("w_bf16", ctypes.c_void_p),
        ("recv", ctypes.c_void_p * MAX_RANKS), ("wpeer", ctypes.c_void_p * MAX_RANKS),
        ("wbf_peer", ctypes.c_void_p * MAX_RANKS), ("flags", ctypes.c_void_p * MAX_RANKS),
        ("epoch", ctypes.c_void_p), ("metrics", ctypes.c_void_p), ("hp", ctypes.c_void_p),
        ("done_counter", ctypes.c_void_p), ("status", ctypes.c_void_p),
        ("timeout_ns", ctypes.c_ulonglong),
        ("pp_lo", ctypes.c_long), ("pp_hi", ctypes.c_long), ("pp_done", ctypes.c_void_p),
        ("mc_g", ctypes.c_void_p), ("mc_w", ctypes.c_void_p), ("mc_flags", ctypes.c_void_p),
        ("nv_flags", ctypes.c_void_p),
    ]


class _PushDesc(ctypes.Structure):
    """TfdePush of csrc/kernels/gemm_fp32.cu: lets a filter-gradient GEMM store its tiles straight into the
    peers' receive slots (GEMM -> reduce-scatter in one kernel)."""
    _fields_ = [
        ("enabled", ctypes.c_int), ("world", ctypes.c_int), ("rank", ctypes.c_int), ("mode", ctypes.c_int),
        ("slot_floats", ctypes.c_long), ("arena_off", ctypes.c_long),
        ("recv", ctypes.c_void_p * MAX_RANKS), ("epoch", ctypes.c_void_p), ("done_flag", ctypes.c_void_p),
    ]


for _n, _a, _r in [
    ("tfde_set_device", [ctypes.c_int], ctypes.c_int),
    ("tfde_symm_alloc", [ctypes.c_size_t, ctypes.POINTER(ctypes.c_void_p)], ctypes.c_int),
    ("tfde_symm_free", [ctypes.c_void_p], ctypes.c_int),
    ("tfde_ipc_get_handle", [ctypes.c_void_p, ctypes.c_void_p], ctypes.c_int),
    ("tfde_ipc_open_handle", [ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p)], ctypes.c_int),
    ("tfde_ipc_close_handle", [ctypes.c_void_p], ctypes.c_int),
    ("tfde_ipc_handle_bytes", [], ctypes.c_int),
    ("tfde_ar_desc_bytes", [], ctypes.c_size_t),
    ("tfde_ar_flag_bytes", [ctypes.c_int], ctypes.c_size_t),
    ("tfde_ar_nvls_flag_bytes", [], ctypes.c_size_t),
    ("tfde_allreduce_fused", [ctypes.POINTER(_ARDesc), ctypes.c_void_p], ctypes.c_int),
    ("tfde_p2p_copy", [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_long, ctypes.c_void_p], ctypes.c_int),
    ("tfde_dense_wgrad_push", [ctypes.c_void_p] * 5 + [ctypes.c_int] * 3 + [ctypes.c_void_p, ctypes.c_size_t,
                               ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p], ctypes.c_int),
    ("tfde_push_desc_bytes", [], ctypes.c_size_t),
]:
    native.register_optional(_n, _a, _r)


class _RawCuda:
    """Minimal __cuda_array_interface__ holder so torch can view memory we cudaMalloc'ed ourselves."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False),
                                         "version": 2, "strides": None}


class SymmBuffer:
    """One symmetric allocation: same size on every rank, every rank holds a mapping of every peer's."""

    def __init__(self, comm: "P2PComm", nbytes: int, tag: str):
        L = native.load()
        self.comm, self.nbytes, self.tag = comm, nbytes, tag
        p = ctypes.c_void_p()
        native.check(L.tfde_symm_alloc(nbytes, ctypes.byref(p)), f"symm_alloc({tag})")
        self.local_ptr = p.value
        self.peer_ptrs: List[int] = [0] * comm.world
        self.peer_ptrs[comm.rank] = self.local_ptr
        self._opened = []
        if comm.world > 1:
            hb = L.tfde_ipc_handle_bytes()
            hbuf = ctypes.create_string_buffer(hb)
            native.check(L.tfde_ipc_get_handle(self.local_ptr, hbuf), "ipc_get_handle")
            handles = [None] * comm.world
            dist.all_gather_object(handles, bytes(hbuf.raw), group=comm.strategy.group)
            for r, h in enumerate(handles):
                if r == comm.rank:
                    continue
                q = ctypes.c_void_p()
                native.check(L.tfde_ipc_open_handle(ctypes.create_string_buffer(h, hb), ctypes.byref(q)),
                             f"ipc_open_handle({tag}, peer {r})")
                self.peer_ptrs[r] = q.value
                self._opened.append(q.value)
        self._holder = _RawCuda(self.local_ptr, nbytes)
        self.bytes_tensor = torch.as_tensor(self._holder, device=comm.device)

    def view(self, dtype, numel=None):
        t = self.bytes_tensor.view(dtype)
        return t if numel is None else t[:numel]

    def free(self):
        L = native.load()
        for q in self._opened:
            L.tfde_ipc_close_handle(q)
        self._opened = []
        if self.local_ptr:
            L.tfde_symm_free(self.local_ptr)
            self.local_ptr = 0


class McBuffer:
    """Symmetric allocation that is ALSO bound to an NVSwitch multicast object (NVLS): ``mc_ptr`` is one
    address that reaches all replicas -- ``multimem.ld_reduce`` through it returns the sum over ranks,
    ``multimem.st`` writes every replica. The VMM allocation, the file-descriptor exchange and the multicast
    bind are torch's symmetric-memory plumbing (set-up only); every per-step access is our kernels'."""

    def __init__(self, comm: "P2PComm", nbytes: int, tag: str):
        from torch._C._distributed_c10d import _SymmetricMemory
        self.comm, self.nbytes, self.tag = comm, nbytes, tag
        group = comm.strategy.group if comm.strategy.group is not None else dist.group.WORLD
        self.bytes_tensor = _SymmetricMemory.empty_strided_p2p((nbytes,), (1,), torch.uint8, comm.device,
                                                               group.group_name)
        self.bytes_tensor.zero_()
        torch.cuda.synchronize(comm.device)
        self.handle = _SymmetricMemory.rendezvous(self.bytes_tensor, group.group_name)
        off = int(getattr(self.handle, "offset", 0) or 0)
        self.local_ptr = self.bytes_tensor.data_ptr()
        self.peer_ptrs = [int(q) + off for q in self.handle.buffer_ptrs]
        self.peer_ptrs[comm.rank] = self.local_ptr
        mc = int(self.handle.multicast_ptr or 0)
        self.mc_ptr = mc + off if mc else 0

    def view(self, dtype, numel=None):
        t = self.bytes_tensor.view(dtype)
        return t if numel is None else t[:numel]

    def free(self):
        self.handle = None
        self.bytes_tensor = None
        self.local_ptr = 0


class P2PComm(CommBase):
    """Symmetric heap + fused all-reduce. One instance per strategy; one "channel" (flags, epochs,
    receive slots) per arena so that independent collectives never share signal state."""

    def __init__(self, strategy):
        super().__init__(strategy)
        if self.world > MAX_RANKS:
            raise ValueError(f"P2PComm supports up to {MAX_RANKS} ranks on one NVSwitch domain")
        L = native.load()
        native.check(L.tfde_set_device(self.device.index or 0), "set_device")
        self._buffers: List[SymmBuffer] = []
        self._channels = {}
        self._named = {}
        self.timeout_s = float(os.environ.get("TFDE_P2P_TIMEOUT_S", "20"))
        self.one_shot_max_bytes = int(os.environ.get("TFDE_ONE_SHOT_MAX_BYTES", str(4 << 20)))
        self.wire_bf16 = os.environ.get("TFDE_WIRE_BF16", "0") == "1"
        self.nvls_min_bytes = int(os.environ.get("TFDE_NVLS_MIN_BYTES", str(64 << 10)))
        self.nvls_min_world = int(os.environ.get("TFDE_NVLS_MIN_WORLD", "4"))
        self._scalar_channel = None
        self.nvls = self._probe_nvls()

    # ---- symmetric memory --------------------------------------------------------------------
    def _probe_nvls(self) -> bool:
        """NVLS needs multicast objects (NVSwitch + driver support). TFDE_NVLS=0 disables, =1 requires it,
        default: use it when the probe allocation succeeds on EVERY rank."""
        want = os.environ.get("TFDE_NVLS", "auto")
        if self.world < 2 or want == "0":
            return False
        ok = 0
        try:
            probe = McBuffer(self, 4096, "nvls_probe")
            ok = 1 if probe.mc_ptr else 0
            self._buffers.append(probe)
        except Exception as ex:  # noqa: BLE001
            log.info("NVLS probe failed on rank %d: %s", self.rank, ex)
        flags = [None] * self.world
        dist.all_gather_object(flags, ok, group=self.strategy.group)
        good = all(flags)
        if want == "1" and not good:
            raise RuntimeError(f"TFDE_NVLS=1 but multicast memory is unavailable (per-rank probe: {flags})")
        log.info("NVLS multicast %s", "available" if good else f"unavailable (per-rank probe: {flags})")
        return good

    def symm_buffer(self, nbytes: int, tag: str, multicast: bool = False):
        nbytes = (nbytes + 255) // 256 * 256
        b = McBuffer(self, nbytes, tag) if (multicast and self.nvls) else SymmBuffer(self, nbytes, tag)
        self._buffers.append(b)
        return b

    def symm_empty(self, numel: int, tag: str, dtype=torch.float32, multicast: bool = False) -> torch.Tensor:
        b = self.symm_buffer(numel * torch.empty((), dtype=dtype).element_size(), tag, multicast)
        t = b.view(dtype, numel)
        self._named[t.data_ptr()] = b
        return t

    def buffer_of(self, t: torch.Tensor) -> Optional[SymmBuffer]:
        return self._named.get(t.data_ptr())

    # ---- small helpers -----------------------------------------------------------------------
    def broadcast_(self, t, src=0):
        """Chief -> all (G13). Symmetric tensors are pulled straight from the chief's memory with the
        p2p copy kernel; anything else goes through a temporary symmetric staging buffer."""
        if self.world == 1:
            return t
        L = native.load()
        buf = self.buffer_of(t)
        flat = t.reshape(-1)
        if buf is None or flat.dtype != torch.float32:
            stage = self.symm_empty(flat.numel(), "bcast_stage", dtype=torch.float32)
            if self.rank == src:
                stage.copy_(flat.to(torch.float32))
            buf = self.buffer_of(stage)
            target = stage
        else:
            target = flat
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.strategy.group)
        if self.rank != src:
            native.check(L.tfde_p2p_copy(target.data_ptr(), buf.peer_ptrs[src], target.numel(),
                                         native.stream_ptr()), "p2p_copy")
            torch.cuda.synchronize(self.device)
        dist.barrier(group=self.strategy.group)
        if target is not flat:
            flat.copy_(target.to(flat.dtype))
        return t

    def all_reduce_sum_(self, t):
        """Small SUM all-reduce (<= 8192 floats) through a dedicated reduce-only channel (G14)."""
        if self.world == 1:
            return t
        flat = t.reshape(-1).to(torch.float32)
        cap = 8192
        if flat.numel() > cap:
            raise ValueError("all_reduce_sum_ is for small tensors; gradients go through reduce_and_apply")
        if self._scalar_channel is None:
            self._scalar_channel = _Channel(self, cap, with_weights=False)
        ch = self._scalar_channel
        ch.g.zero_()
        ch.g[:flat.numel()].copy_(flat)
        ch.launch(opt=-1, hp=None, s1=None, s2=None)
        t.copy_(ch.w[:flat.numel()].reshape(t.shape).to(t.dtype))
        return t

    # ---- the hot path ------------------------------------------------------------------------
    def prepare(self, arena: Arena, optimizer):
        # the channel lives ON the arena (an id()-keyed dict would hand a recycled id the stale channel
        # of a dead arena on one rank only -> ranks disagree about epochs -> dead-lock)
        ch = getattr(arena, "_channel", None)
        if ch is None or ch.comm is not self:
            ch = _Channel(self, arena.n_total, with_weights=True, arena=arena)
            arena._channel = ch
            self._channels[len(self._channels)] = ch
        ch.configure_optimizer(optimizer, arena)
        return ch

    def reduce_and_apply(self, arena: Arena, optimizer):
        ch = getattr(arena, "_channel", None)
        if ch is None or ch.comm is not self:
            ch = self.prepare(arena, optimizer)
        ch.launch(opt=optimizer.native_code, hp=optimizer.hparams, s1=arena.s1, s2=arena.s2)

    def check_health(self):
        for ch in list(self._channels.values()) + ([self._scalar_channel] if self._scalar_channel else []):
            ch.check()

    def close(self):
        try:
            torch.cuda.synchronize(self.device)
        except Exception:
            pass
        for b in self._buffers:
            try:
                b.free()
            except Exception:
                pass
        self._buffers = []
        self._channels = {}
        from ..ops import push as _push
        _push.clear()      # descriptors point into the buffers just freed


class _Channel:
    """Signal pad + receive slots + launch descriptor for one arena."""

    def __init__(self, comm: P2PComm, n_total: int, with_weights: bool, arena: Optional[Arena] = None):
        L = native.load()
        self.comm = comm
        W = comm.world
        self.n_total = n_total
        bytes_total = n_total * 4
        self.mode = ONE_SHOT if (W <= 2 or bytes_total * (W - 1) <= comm.one_shot_max_bytes) else TWO_SHOT
        # NVLS needs both arenas multicast-bound (Arena does that when comm.nvls) and fp32 on the wire
        g_mc = comm.buffer_of(arena.g) if arena is not None else None
        w_mc = comm.buffer_of(arena.w) if arena is not None else None
        nvls_ok = (comm.nvls and W >= 2 and not comm.wire_bf16 and getattr(g_mc, "mc_ptr", 0)
                   and getattr(w_mc, "mc_ptr", 0))
        # measured (profiles/allreduce_sweep_n{2,8}): at 8 GPUs the switch-reduced variant wins from 256 KB up
        # (15 vs 23 us at Model A's 1.39 MB, 727 vs 559 GB/s bus at 256 MB); at 2 GPUs it moves 1.5x the
        # bytes of two-shot and loses everywhere
        if nvls_ok and W >= comm.nvls_min_world and bytes_total >= comm.nvls_min_bytes:
            self.mode = NVLS
        forced = os.environ.get("TFDE_AR_MODE")
        if forced in ("one_shot", "two_shot"):
            self.mode = ONE_SHOT if forced == "one_shot" else TWO_SHOT
        elif forced == "nvls":
            if not nvls_ok:
                raise RuntimeError("TFDE_AR_MODE=nvls but NVLS multicast memory is not available for this arena")
            self.mode = NVLS
        if self.mode == ONE_SHOT:
            self.slot = n_total
        else:
            self.slot = ((n_total + W - 1) // W + 3) // 4 * 4
        if self.mode == NVLS:
            self.recv = None
            self.flags = comm.symm_buffer(L.tfde_ar_nvls_flag_bytes(), "nvls_flags", multicast=True)
        else:
            self.recv = comm.symm_buffer(2 * W * self.slot * 4, "recv")
            self.flags = comm.symm_buffer(L.tfde_ar_flag_bytes(W), "flags")
        dev = comm.device
        self.epoch = torch.zeros(MAX_BLOCKS, dtype=torch.int32, device=dev)
        self.done = torch.zeros(1, dtype=torch.int32, device=dev)
        self.status = torch.zeros(4, dtype=torch.int32, device=dev)
        if arena is not None:
            self.g, self.w, self.metrics = arena.g, arena.w, arena.metrics
            self.n, self.n_extra = arena.n, arena.n_extra
            self.wbuf = comm.buffer_of(arena.w)
            if self.wbuf is None and W > 1:
                raise RuntimeError("parameter arena must live in symmetric memory (Arena(..., comm=...))")
        else:
            self.g = torch.zeros(n_total, dtype=torch.float32, device=dev)
            self.w = comm.symm_empty(n_total, "scalars_w")
            self.wbuf = comm.buffer_of(self.w)
            self.metrics = None
            self.n, self.n_extra = n_total, 0
        default_blocks = 16 if bytes_total <= (256 << 10) else (64 if bytes_total <= (4 << 20) else 144)
        if self.mode == NVLS:      # the switch does the fan-in/fan-out: few blocks saturate it
            default_blocks = 8 if bytes_total <= (4 << 20) else 32      # 32 > 64/128 at every size measured
        self.blocks = int(os.environ.get("TFDE_AR_BLOCKS", str(default_blocks)))
        self.blocks = max(1, min(self.blocks, MAX_BLOCKS))
        self.desc = _ARDesc()
        d = self.desc
        d.rank, d.world, d.mode = comm.rank, W, self.mode
        d.wire_bf16 = 1 if comm.wire_bf16 else 0
        d.n_extra, d.blocks = self.n_extra, self.blocks
        d.n, d.n_total, d.slot_floats = self.n, n_total, self.slot
        d.g, d.w = self.g.data_ptr(), self.w.data_ptr()
        for r in range(W):
            d.recv[r] = self.recv.peer_ptrs[r] if self.recv is not None else None
            d.flags[r] = self.flags.peer_ptrs[r]
            d.wpeer[r] = self.wbuf.peer_ptrs[r] if self.wbuf is not None else self.w.data_ptr()
        if self.mode == NVLS:
            d.mc_g, d.mc_w = g_mc.mc_ptr, w_mc.mc_ptr
            d.mc_flags, d.nv_flags = self.flags.mc_ptr, self.flags.local_ptr
        d.epoch, d.done_counter, d.status = self.epoch.data_ptr(), self.done.data_ptr(), self.status.data_ptr()
        d.metrics = self.metrics.data_ptr() if self.metrics is not None else None
        d.timeout_ns = int(comm.timeout_s * 1e9)
        # all ranks must have mapped everything before anyone launches
        if W > 1:
            torch.cuda.synchronize(dev)
            dist.barrier(group=comm.strategy.group)
        log.info("p2p channel: n_total=%d mode=%s slot=%d blocks=%d wire_bf16=%s", n_total,
                 _MODE_NAMES[self.mode], self.slot, self.blocks, comm.wire_bf16)

    def enable_fused_push(self, grad_view: torch.Tensor, arena_offset: int):
        """Registers one gradient tensor (the largest dense kernel) whose wgrad GEMM pushes its tiles into
        the peers' receive slots itself; the all-reduce kernel then skips that range when the GEMM says it
        ran this step (pp_done == epoch)."""
        if (self.comm.world < 2 or self.comm.wire_bf16 or self.mode == NVLS
                or os.environ.get("TFDE_FUSED_PUSH", "1") != "1"):
            return False      # NVLS has no receive slots to push into: the switch reads the gradients in place
        n = grad_view.numel()
        if n % 4 or arena_offset % 4 or ctypes.sizeof(_PushDesc) != native.load().tfde_push_desc_bytes():
            return False
        self.pp_done = torch.zeros(1, dtype=torch.int32, device=self.comm.device)
        pd = _PushDesc()
        pd.enabled, pd.world, pd.rank, pd.mode = 1, self.comm.world, self.comm.rank, self.mode
        pd.slot_floats, pd.arena_off = self.slot, arena_offset
        for r in range(self.comm.world):
            pd.recv[r] = self.recv.peer_ptrs[r]
        pd.epoch, pd.done_flag = self.epoch.data_ptr(), self.pp_done.data_ptr()
        self.push_desc = pd
        self.desc.pp_lo, self.desc.pp_hi, self.desc.pp_done = arena_offset, arena_offset + n, self.pp_done.data_ptr()
        native.register_wgrad_push(grad_view.data_ptr(), pd)
        log.info("fused wgrad->push enabled for arena range [%d, %d)", arena_offset, arena_offset + n)
        return True

    def configure_optimizer(self, optimizer, arena: Arena):
        """Allocates optimizer slots with the layout this channel's mode needs (two-shot: only the owned
        slice -- optimizer state is sharded by owner rank)."""
        n_state = self.slot if self.mode in (TWO_SHOT, NVLS) else arena.n_total
        need = optimizer.num_slots
        if need >= 1 and (arena.s1 is None or arena.s1.numel() != n_state):
            arena.s1 = torch.zeros(n_state, dtype=torch.float32, device=arena.device)
        if need >= 2 and (arena.s2 is None or arena.s2.numel() != n_state):
            arena.s2 = torch.zeros(n_state, dtype=torch.float32, device=arena.device)
        arena.state_sharded = self.mode in (TWO_SHOT, NVLS) and self.comm.world > 1
        arena.state_slice = (self.comm.rank * self.slot, self.slot)

    def launch(self, opt, hp, s1, s2):
        L = native.load()
        d = self.desc
        d.opt = opt
        d.hp = hp.data_ptr() if hp is not None else None
        d.s1 = s1.data_ptr() if s1 is not None else None
        d.s2 = s2.data_ptr() if s2 is not None else None
        native.check(L.tfde_allreduce_fused(ctypes.byref(d), native.stream_ptr()), "allreduce_fused")

    def check(self):
        st = self.status.cpu().tolist()
        if st[0] != 0:
            raise RuntimeError(
                f"p2p all-reduce timed out on rank {self.comm.rank}: epoch {st[1]}, waiting for peer "
                f"{st[2]}, block {st[3]} (peer dead or diverged; restart from the last checkpoint)")
