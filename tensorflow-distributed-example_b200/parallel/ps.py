"""Asynchronous data parallelism with parameter servers (S2 / N4 / G16 in SURVEY.md).

Reference: ``tf.contrib.distribute.ParameterServerStrategy()`` as ``train_distribute``
(mnist_keras_distributed.py:241-242) and ``tf.distribute.experimental.ParameterServerStrategy()``
(tf2_mnist_distributed.py:189), with ``ps`` / ``worker`` / ``master`` roles taken from TF_CONFIG and
per-role device filters (mnist_keras_distributed.py:165-189).

Semantics reproduced: variables are placed round-robin on the ``ps`` tasks; every worker, on its own
clock, pulls the current weights, computes gradients on ITS batch (loss normalised by the per-worker
batch -- there is no global batch in async training), and pushes them; the owner applies the update
immediately, without waiting for other workers (Hogwild-style); ``global_step`` lives on ps 0 and
counts applied updates from all workers; ``ps`` processes only serve; a dead worker does not stop the
others; only the chief writes checkpoints.

Two transports, chosen per ps task when the chief initialises it:

* **nvlink** (B200-native, default when the ps task and the workers are GPU processes on one host, i.e. one
  NVLink/NVSwitch domain, and the optimizer is plain SGD -- the reference's case): the ps task keeps its variables
  in ONE flat shard of its GPU's memory and exports it with CUDA IPC. A worker maps the shard once; per step it
  *pulls* with one kernel of plain peer loads (``tfde_ps_pull_multi``) and *pushes* with one kernel of NVLink
  reduction atomics, ``owner -= lr * g`` (``tfde_ps_push_multi``), which also bumps ``global_step`` -- a device
  counter on ps 0. No host round trip, no serialisation, no lock: updates from different workers interleave exactly
  like TF's unsynchronised async SGD. The TCP channel below carries only control messages (init, attach, done).
* **tcp** (cross-host fallback, CPU-only boxes, optimizers with slots): each ps task listens on ITS ``host:port``
  from the cluster spec (what TF's gRPC server does); workers connect only to the ps tasks (the device-filter rule).
  Messages are a JSON header plus raw array bytes -- nothing on the wire is ever unpickled -- behind an HMAC
  handshake whose key comes from ``TFDE_PS_AUTHKEY`` / ``TFDE_PS_AUTHKEY_FILE``; the built-in default key is only
  accepted for loopback listeners.

With no TF_CONFIG the strategy degrades to one local replica, like TF's local mode.
"""
from __future__ import annotations

import ctypes
import json
import logging
import os
import struct
import threading
import time
from multiprocessing.connection import Client, Listener
from typing import Dict, List, Optional

import numpy as np
import torch

from ..cluster import tf_config as tfc
from .strategy import Strategy, _pick_device

log = logging.getLogger("tfde_b200")
_DEFAULT_KEY = b"tfde_b200-ps"
_LOOPBACK = ("127.0.0.1", "localhost", "::1")
_DTYPES = {"float32": np.float32, "float64": np.float64, "int32": np.int32, "int64": np.int64, "uint8": np.uint8}
_MAX_MSG = 1 << 31


def _authkey() -> bytes:
    k = os.environ.get("TFDE_PS_AUTHKEY")
    if k:
        return k.encode()
    f = os.environ.get("TFDE_PS_AUTHKEY_FILE")
    if f:
        with open(f, "rb") as fh:
            return fh.read().strip()
    return _DEFAULT_KEY


# ----------------------------------------------------------------------------------------------
# wire format: <u32 header length><JSON header><raw array bytes...>  (no pickle anywhere)
# ----------------------------------------------------------------------------------------------
def _send(conn, op: str, meta: Optional[dict] = None, arrays: Optional[Dict[str, np.ndarray]] = None):
    arrays = arrays or {}
    blobs, desc = [], []
    for k, a in arrays.items():
        a = np.ascontiguousarray(a)
        if a.dtype.name not in _DTYPES:
            raise TypeError(f"{k}: dtype {a.dtype} is not allowed on the ps wire")
        desc.append([k, a.dtype.name, list(a.shape)])
        blobs.append(a.tobytes())
    hb = json.dumps({"op": op, "meta": meta or {}, "arrays": desc}).encode()
    conn.send_bytes(struct.pack("<I", len(hb)) + hb + b"".join(blobs))


def _recv(conn):
    buf = conn.recv_bytes(_MAX_MSG)
    if len(buf) < 4:
        raise ValueError("short ps message")
    n, = struct.unpack_from("<I", buf)
    if 4 + n > len(buf):
        raise ValueError("corrupt ps message header")
    hdr = json.loads(buf[4:4 + n].decode())
    off = 4 + n
    arrays = {}
    for name, dt, shape in hdr.get("arrays", []):
        if dt not in _DTYPES:
            raise ValueError(f"dtype {dt!r} not allowed on the ps wire")
        shape = tuple(int(s) for s in shape)
        cnt = int(np.prod(shape)) if shape else 1
        nbytes = cnt * np.dtype(_DTYPES[dt]).itemsize
        if cnt < 0 or off + nbytes > len(buf):
            raise ValueError("corrupt ps message payload")
        arrays[str(name)] = np.frombuffer(buf, dtype=_DTYPES[dt], count=cnt, offset=off).reshape(shape).copy()
        off += nbytes
    return str(hdr.get("op")), dict(hdr.get("meta") or {}), arrays


class _RawCuda:
    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False),
                                         "version": 2, "strides": None}


def _align4(n: int) -> int:
    return (n + 3) // 4 * 4


class _ServerState:
    """Variables of one ps task: one flat fp32 shard (host memory, or -- nvlink transport -- device memory exported
    over CUDA IPC) + optimizer slots + the global step (ps 0)."""

    def __init__(self, index: int):
        self.index = index
        self.device = torch.device("cpu")
        self.flat: Optional[torch.Tensor] = None
        self.step_t: Optional[torch.Tensor] = None          # int32[1] view of the shard's last word (nvlink)
        self.layout: Dict[str, tuple] = {}                  # name -> (offset, size, shape, trainable)
        self.vars: Dict[str, torch.Tensor] = {}
        self.slots: Dict[str, List[torch.Tensor]] = {}
        self.count: Dict[str, int] = {}                     # updates applied per variable (Adam bias correction)
        self.opt = {"kind": "sgd", "momentum": 0.0, "nesterov": False, "beta_1": 0.9, "beta_2": 0.999,
                    "epsilon": 1e-7}
        self._step = 0
        self.base_step = 0
        self.transport = "tcp"
        self.raw_ptr = 0
        self.nbytes = 0
        self.initialized = False
        self.done = set()
        self.shutdown = threading.Event()
        self.lock = threading.Lock()

    # ---- global step -------------------------------------------------------------------------
    @property
    def global_step(self) -> int:
        if self.step_t is not None:
            return int(self.step_t.item())
        return self._step

    def set_step(self, v: int):
        if self.step_t is not None:
            self.step_t.fill_(int(v))
        self._step = int(v)

    # ---- set-up ------------------------------------------------------------------------------
    def init(self, arrays: Dict[str, np.ndarray], meta: dict):
        trainable = meta.get("trainable", {})
        want = meta.get("transport", "tcp")
        self.opt.update(meta.get("opt") or {})
        off = 0
        for k, a in arrays.items():
            self.layout[k] = (off, int(a.size), tuple(a.shape), bool(trainable.get(k, True)))
            off += _align4(int(a.size))
        total = off + 4
        self.transport = "tcp"
        if want == "nvlink" and self.opt["kind"] == "sgd" and torch.cuda.is_available():
            try:
                self._alloc_device(total, meta)
                self.transport = "nvlink"
            except Exception as ex:  # noqa: BLE001  (no usable GPU on this ps task: serve over TCP)
                log.warning("ps %d: nvlink transport unavailable (%s); serving over TCP", self.index, ex)
        if self.transport == "tcp":
            self.flat = torch.zeros(total, dtype=torch.float32)
        for k, a in arrays.items():
            o, n, shape, _ = self.layout[k]
            self.vars[k] = self.flat[o:o + n].view(shape)
            self.vars[k].copy_(torch.from_numpy(np.asarray(a, dtype=np.float32)).to(self.device))
            self.count[k] = 0
        self.base_step = int(meta.get("step", 0))
        self.set_step(self.base_step)
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
        self.initialized = True

    def _alloc_device(self, total_floats: int, meta: dict):
        from ..ops import native
        L = native.load()
        ndev = torch.cuda.device_count()
        idx = os.environ.get("TFDE_PS_DEVICE")
        idx = int(idx) if idx not in (None, "") else (int(meta.get("n_workers", 0)) + self.index) % max(ndev, 1)
        self.device = torch.device("cuda", idx % max(ndev, 1))
        torch.cuda.set_device(self.device)
        for name, args, res in [("tfde_symm_alloc", [ctypes.c_size_t, ctypes.POINTER(ctypes.c_void_p)], ctypes.c_int),
                                ("tfde_ipc_get_handle", [ctypes.c_void_p, ctypes.c_void_p], ctypes.c_int),
                                ("tfde_ipc_handle_bytes", [], ctypes.c_int)]:
            native.register_optional(name, args, res)
        self.nbytes = (total_floats * 4 + 255) // 256 * 256
        p = ctypes.c_void_p()
        native.check(L.tfde_symm_alloc(self.nbytes, ctypes.byref(p)), "ps shard alloc", kernels=0)
        self.raw_ptr = p.value
        hb = L.tfde_ipc_handle_bytes()
        hbuf = ctypes.create_string_buffer(hb)
        native.check(L.tfde_ipc_get_handle(self.raw_ptr, hbuf), "ps shard ipc handle", kernels=0)
        self.handle_hex = bytes(hbuf.raw).hex()
        self._holder = _RawCuda(self.raw_ptr, self.nbytes)
        raw = torch.as_tensor(self._holder, device=self.device)
        self.flat = raw.view(torch.float32)[:total_floats]
        self.step_off = total_floats - 1
        self.step_t = raw.view(torch.int32)[self.step_off:self.step_off + 1]

    def attach_info(self) -> dict:
        return {"transport": self.transport, "handle": getattr(self, "handle_hex", ""),
                "device": self.device.index if self.device.type == "cuda" else -1, "nbytes": self.nbytes,
                "layout": {k: [v[0], v[1], 1 if v[3] else 0] for k, v in self.layout.items()},
                "step_off": getattr(self, "step_off", -1)}

    # ---- TCP-side update -----------------------------------------------------------------------
    def apply(self, name: str, g: torch.Tensor, lr: float):
        w = self.vars[name]
        g = g.to(w.device).view_as(w)
        kind = self.opt["kind"]
        self.count[name] = self.count.get(name, 0) + 1
        if kind == "sgd":
            w.sub_(g, alpha=lr)                  # in place; concurrent workers interleave (async SGD)
        elif kind in ("momentum", "nesterov"):
            v = self.slots.setdefault(name, [torch.zeros_like(w)])[0]
            mu = self.opt["momentum"]
            v.mul_(mu).sub_(g, alpha=lr)
            w.add_(mu * v - lr * g if kind == "nesterov" else v)
        else:
            # Adam. The bias correction uses THIS variable's own update count: global_step only advances on ps 0,
            # so variables placed on other ps tasks would otherwise stay at t = 1 forever.
            sl = self.slots.setdefault(name, [torch.zeros_like(w), torch.zeros_like(w)])
            b1, b2, eps = self.opt["beta_1"], self.opt["beta_2"], self.opt["epsilon"]
            t = self.base_step + self.count[name]
            sl[0].mul_(b1).add_(g, alpha=1 - b1)
            sl[1].mul_(b2).addcmul_(g, g, value=1 - b2)
            w.sub_(sl[0] / (sl[1].sqrt() + eps), alpha=lr * float(np.sqrt(1 - b2 ** t) / (1 - b1 ** t)))

    def free(self):
        if self.raw_ptr:
            try:
                from ..ops import native
                torch.cuda.synchronize(self.device)
                self.vars, self.flat, self.step_t = {}, None, None
                native.load().tfde_symm_free(self.raw_ptr)
            except Exception:  # noqa: BLE001
                pass
            self.raw_ptr = 0


def _serve_connection(conn, st: _ServerState, counts_steps: bool):
    try:
        while not st.shutdown.is_set():
            try:
                op, meta, arrays = _recv(conn)
            except (EOFError, ConnectionResetError, OSError):
                return
            except ValueError as ex:
                _send(conn, "error", {"why": str(ex)})
                return
            if op == "init":
                with st.lock:
                    fresh = not st.initialized
                    if fresh:
                        st.init(arrays, meta)
                _send(conn, "ok", {"fresh": fresh, **st.attach_info()})
            elif op == "wait_init":
                t0 = time.time()
                while not st.initialized and time.time() - t0 < float(meta.get("timeout", 60)) and not st.shutdown.is_set():
                    time.sleep(0.01)
                _send(conn, "ok", {"initialized": st.initialized, **(st.attach_info() if st.initialized else {})})
            elif op == "pull":
                names = [k for k in meta.get("names", []) if k in st.vars]
                _send(conn, "ok", {"step": st.global_step}, {k: st.vars[k].detach().cpu().numpy() for k in names})
            elif op == "push":
                lr = float(meta.get("lr", 0.0))
                for k, a in arrays.items():
                    kind, name = k[:2], k[2:]
                    if name not in st.vars:
                        continue
                    if kind == "g:":
                        st.apply(name, torch.from_numpy(a), lr)
                    elif kind == "a:":
                        st.vars[name].copy_(torch.from_numpy(a).to(st.vars[name].device).view_as(st.vars[name]))
                if counts_steps:
                    with st.lock:
                        st.set_step(st.global_step + 1)
                _send(conn, "ok", {"step": st.global_step})
            elif op == "step":
                _send(conn, "ok", {"step": st.global_step})
            elif op == "set_step":
                with st.lock:
                    st.set_step(max(st.global_step, int(meta.get("step", 0))))
                _send(conn, "ok", {"step": st.global_step})
            elif op == "done":
                with st.lock:
                    st.done.add(str(meta.get("worker")))
                _send(conn, "ok", {"n": len(st.done)})
            elif op == "num_done":
                _send(conn, "ok", {"n": len(st.done)})
            elif op == "shutdown":
                st.shutdown.set()
                _send(conn, "ok", {})
                return
            else:
                _send(conn, "error", {"why": f"unknown op {op!r}"})
    finally:
        try:
            conn.close()
        except Exception:  # noqa: BLE001
            pass


class ParameterServer:
    """One ``ps`` task. ``serve_forever`` returns after a ``shutdown`` message (sent by the chief when
    training ends) -- TF's ps tasks never return; a clean exit is friendlier to schedulers."""

    def __init__(self, host: str, port: int, index: int):
        self.host, self.port, self.index = host, port, index
        self.state = _ServerState(index)
        self._listener = None

    def serve_forever(self):
        key = _authkey()
        if key == _DEFAULT_KEY and self.host not in _LOOPBACK:
            raise RuntimeError(
                f"ps {self.index}: refusing to listen on {self.host}:{self.port} with the built-in default key; set "
                "TFDE_PS_AUTHKEY (or TFDE_PS_AUTHKEY_FILE) to a per-job secret on every task of the cluster")
        self._listener = Listener((self.host, self.port), authkey=key)
        log.info("ps %d serving on %s:%d", self.index, self.host, self.port)
        threads = []

        def acceptor():
            while not self.state.shutdown.is_set():
                try:
                    conn = self._listener.accept()
                except (OSError, EOFError):
                    return
                except Exception:   # noqa: BLE001  (failed handshake from a stray client)
                    continue
                t = threading.Thread(target=_serve_connection, args=(conn, self.state, self.index == 0),
                                     daemon=True)
                t.start()
                threads.append(t)

        acc = threading.Thread(target=acceptor, daemon=True)
        acc.start()
        self.state.shutdown.wait()
        time.sleep(0.05)
        try:
            self._listener.close()
        except Exception:  # noqa: BLE001
            pass
        log.info("ps %d shut down at global_step %d (%s transport)", self.index, self.state.global_step,
                 self.state.transport)
        self.state.free()


class _Shard:
    """A worker's mapping of one ps task's device shard (nvlink transport)."""

    def __init__(self):
        self.ptr = 0                 # peer-mapped base of the owner's shard
        self.pull_table = None       # device tensors of TfdePsSeg records
        self.push_table = None
        self.n_pull = self.n_push = 0
        self.max_n = 0
        self.step_ptr = 0


class ParameterServerStrategy(Strategy):
    """``ParameterServerStrategy()``: no-arg constructor that reads TF_CONFIG."""

    def __init__(self, cluster_resolver: Optional[tfc.ClusterResolver] = None, device: Optional[str] = None):
        self.resolver = cluster_resolver or tfc.resolve_from_env()
        r = self.resolver
        self.rank, self.world_size = 0, 1          # every worker is its own replica (async)
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        if r.distributed and r.task.type in tfc.TRAIN_JOBS and device is None and torch.cuda.is_available():
            self.local_rank = r.rank % max(torch.cuda.device_count(), 1)
        self.device = _pick_device(self.local_rank, device) if not r.is_ps else torch.device("cpu")
        if self.device.type == "cuda":
            torch.cuda.set_device(self.device)
        self.group, self._comm = None, None
        self._clients: List = []
        self._owner: Dict[str, int] = {}
        self._shards: Dict[int, _Shard] = {}
        self._step = 0
        self._step_host = None
        self.worker_id = f"{r.task.type}:{r.task.index}"
        self.device_filters = r.device_filters()

    # ---- roles ---------------------------------------------------------------------------------
    @property
    def distributed(self) -> bool:
        return self.resolver.distributed and len(self.resolver.ps_endpoints()) > 0

    @property
    def synchronous(self) -> bool:
        return not self.distributed

    @property
    def is_chief(self) -> bool:
        return self.resolver.is_chief

    @property
    def replica_id(self) -> int:
        return max(self.resolver.rank, 0)       # distinct dropout / shuffle streams per worker

    def run_server(self):
        r = self.resolver
        if not r.is_ps:
            raise RuntimeError("run_server() is for 'ps' tasks")
        host, port = tfc.split_hostport(r.cluster.task_address("ps", r.task.index))
        ParameterServer(host, port, r.task.index).serve_forever()

    # ---- worker side ---------------------------------------------------------------------------
    def _connect(self):
        if self._clients:
            return
        deadline = time.time() + float(os.environ.get("TFDE_RENDEZVOUS_TIMEOUT", "120"))
        key = _authkey()
        for host, port in self.resolver.ps_endpoints():
            while True:
                try:
                    self._clients.append(Client((host, port), authkey=key))
                    break
                except (ConnectionRefusedError, OSError):
                    if time.time() > deadline:
                        raise TimeoutError(f"parameter server {host}:{port} unreachable")
                    time.sleep(0.1)

    def _call(self, ps: int, op: str, meta: Optional[dict] = None, arrays=None):
        c = self._clients[ps]
        _send(c, op, meta, arrays)
        rop, rmeta, rarrays = _recv(c)
        if rop != "ok":
            raise RuntimeError(f"ps {ps}: {rop} {rmeta}")
        return rmeta, rarrays

    def _wanted_transport(self, model) -> str:
        want = os.environ.get("TFDE_PS_TRANSPORT", "auto")
        if want in ("tcp", "nvlink"):
            return want
        me = None
        try:
            me = tfc.split_hostport(self.resolver.cluster.task_address(self.resolver.task.type,
                                                                       self.resolver.task.index))[0]
        except Exception:  # noqa: BLE001
            pass
        hosts = {h for h, _ in self.resolver.ps_endpoints()}
        same_host = all(h in _LOOPBACK or h == me for h in hosts)
        ok = self.device.type == "cuda" and same_host and model.optimizer.kind == "sgd"
        return "nvlink" if ok else "tcp"

    def on_train_begin(self, model):
        if not self.distributed:
            return
        self._connect()
        nps = len(self._clients)
        self._owner = {v.name: i % nps for i, v in enumerate(model.variables)}   # round-robin placement
        model.use_cuda_graph = False    # the step contains pulls/pushes on the host's clock
        opt = model.optimizer
        cfg = {"kind": opt.kind, "momentum": opt.momentum, "nesterov": getattr(opt, "nesterov", False),
               "beta_1": opt.beta_1, "beta_2": opt.beta_2, "epsilon": opt.epsilon}
        infos = []
        if self.is_chief:
            model.sync()
            named = model.named_weights()
            trainable = {v.name: bool(v.trainable) for v in model.variables}
            for ps in range(nps):
                mine = {k: v for k, v in named.items() if self._owner[k] == ps}
                meta, _ = self._call(ps, "init", {"opt": cfg, "step": int(opt.iterations), "trainable": trainable,
                                                  "transport": self._wanted_transport(model),
                                                  "n_workers": len(self.resolver.cluster.training_tasks())}, mine)
                infos.append(meta)
        else:
            for ps in range(nps):
                meta, _ = self._call(ps, "wait_init", {"timeout": float(os.environ.get("TFDE_RENDEZVOUS_TIMEOUT", "120"))})
                if not meta.get("initialized"):
                    raise TimeoutError("chief never initialised the parameter servers")
                infos.append(meta)
        self._shards = {}
        for ps, info in enumerate(infos):
            if info.get("transport") == "nvlink" and self.device.type == "cuda" and \
                    os.environ.get("TFDE_PS_TRANSPORT", "auto") != "tcp":
                self._shards[ps] = self._attach(ps, info, model)
        self.transport = "nvlink" if len(self._shards) == nps else ("mixed" if self._shards else "tcp")
        log.info("parameter-server transport: %s (%d ps task(s))", self.transport, nps)
        self.pull(model)

    def _attach(self, ps: int, info: dict, model) -> _Shard:
        """Maps ps task `ps`'s device shard (CUDA IPC) and builds the segment tables of the pull/push kernels."""
        from ..ops import native
        L = native.load()
        for name, args, res in [
            ("tfde_ipc_open_handle", [ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p)], ctypes.c_int),
            ("tfde_ipc_close_handle", [ctypes.c_void_p], ctypes.c_int),
            ("tfde_ps_seg_bytes", [], ctypes.c_size_t),
            ("tfde_ps_pull_multi", [ctypes.c_void_p, ctypes.c_int, ctypes.c_long, ctypes.c_void_p], ctypes.c_int),
            ("tfde_ps_push_multi", [ctypes.c_void_p, ctypes.c_int, ctypes.c_long, ctypes.c_float, ctypes.c_void_p,
                                    ctypes.c_void_p, ctypes.c_void_p], ctypes.c_int),
        ]:
            native.register_optional(name, args, res)
        if L.tfde_ps_seg_bytes() != 32:
            raise RuntimeError("TfdePsSeg layout mismatch")
        sh = _Shard()
        h = bytes.fromhex(info["handle"])
        q = ctypes.c_void_p()
        native.check(L.tfde_ipc_open_handle(ctypes.create_string_buffer(h, len(h)), ctypes.byref(q)),
                     f"ipc_open_handle(ps {ps})", kernels=0)
        sh.ptr = q.value
        pull, push = [], []
        for v in model.variables:
            if self._owner[v.name] != ps:
                continue
            off, size, _tr = info["layout"][v.name]
            owner = sh.ptr + 4 * int(off)
            pull += [owner, v.value.data_ptr(), int(size), 0]
            if v.trainable:
                push += [owner, v.grad.data_ptr(), int(size), 0]
            else:
                push += [owner, v.value.data_ptr(), int(size), 1]
            sh.max_n = max(sh.max_n, int(size))
        sh.n_pull, sh.n_push = len(pull) // 4, len(push) // 4
        sh.pull_table = torch.tensor(pull or [0], dtype=torch.int64).to(self.device)
        sh.push_table = torch.tensor(push or [0], dtype=torch.int64).to(self.device)
        if ps == 0:
            sh.step_ptr = sh.ptr + 4 * int(info["step_off"])
            if self._step_host is None:
                from ..ops import hostmem
                self._step_host = torch.zeros(1, dtype=torch.int32).pin_memory()
                self._step_dev = hostmem.device_alias(self._step_host, self.device)
        return sh

    def global_step(self) -> int:
        if not self.distributed:
            return self._step
        self._connect()
        meta, _ = self._call(0, "step")
        self._step = int(meta["step"])
        return self._step

    def pull(self, model):
        from ..ops import native
        nps = len(self._clients)
        for ps in range(nps):
            sh = self._shards.get(ps)
            if sh is not None:
                if sh.n_pull:
                    native.check(native.load().tfde_ps_pull_multi(sh.pull_table.data_ptr(), sh.n_pull, sh.max_n,
                                                                  native.stream_ptr()), "ps_pull_multi")
                continue
            names = [v.name for v in model.variables if self._owner[v.name] == ps]
            if not names:
                continue
            meta, arrays = self._call(ps, "pull", {"names": names})
            if ps == 0:
                self._step = int(meta["step"])
            for v in model.variables:
                if v.name in arrays:
                    v.value.copy_(torch.from_numpy(arrays[v.name]).to(v.value.device, non_blocking=True).view_as(v.value))

    def push(self, model):
        from ..ops import native
        lr = model.optimizer.learning_rate
        nps = len(self._clients)
        tcp = [ps for ps in range(nps) if ps not in self._shards]
        for ps, sh in self._shards.items():
            if sh.n_push or ps == 0:
                native.check(native.load().tfde_ps_push_multi(
                    sh.push_table.data_ptr(), sh.n_push, sh.max_n, float(lr),
                    sh.step_ptr or None, self._step_dev.data_ptr() if (ps == 0 and sh.step_ptr) else None,
                    native.stream_ptr()), "ps_push_multi")
        if tcp or 0 in self._shards:
            model.sync()
        if 0 in self._shards:
            self._step = int(self._step_host[0])
        for ps in tcp:
            arrays = {}
            for v in model.variables:
                if self._owner[v.name] != ps:
                    continue
                if v.trainable:
                    arrays["g:" + v.name] = v.grad.detach().cpu().numpy()
                else:
                    arrays["a:" + v.name] = v.value.detach().cpu().numpy()
            meta, _ = self._call(ps, "push", {"lr": float(lr)}, arrays)
            if ps == 0:
                self._step = int(meta["step"])

    def after_step(self, _local_step) -> int:
        return self._step if self.distributed else _local_step + 1

    def on_train_end(self, model):
        if not self.distributed:
            return
        model.sync()
        if not self.is_chief:
            self._detach()          # unmap the ps shards before telling the chief (who releases the ps tasks) we are done
        self._call(0, "done", {"worker": self.worker_id})
        if self.is_chief:
            # wait (bounded) for the other workers, then release the ps tasks
            n_workers = len(self.resolver.cluster.training_tasks())
            deadline = time.time() + float(os.environ.get("TFDE_PS_DRAIN_TIMEOUT", "30"))
            while time.time() < deadline:
                meta, _ = self._call(0, "num_done")
                if int(meta["n"]) >= n_workers:
                    break
                time.sleep(0.05)
            self.pull(model)   # final weights for the last checkpoint / export
            model.sync()
        self._detach()
        if self.is_chief:
            for ps in range(len(self._clients)):
                try:
                    self._call(ps, "shutdown")
                except Exception:  # noqa: BLE001
                    pass
        for c in self._clients:
            try:
                c.close()
            except Exception:  # noqa: BLE001
                pass
        self._clients = []

    def _detach(self):
        if not self._shards:
            return
        from ..ops import native
        L = native.load()
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
        for sh in self._shards.values():
            if sh.ptr:
                L.tfde_ipc_close_handle(sh.ptr)
                sh.ptr = 0
        self._shards = {}
