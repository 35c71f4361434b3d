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

Transport: each ps task listens on ITS ``host:port`` from the cluster spec (what TF's gRPC server
does); workers connect only to the ps tasks (the device-filter rule: a worker talks to ps + itself).
Messages are length-prefixed pickles of numpy arrays over TCP (``multiprocessing.connection``). With no
TF_CONFIG the strategy degrades to one local replica, like TF's local mode.
"""
from __future__ import annotations

import logging
import os
import threading
import time
from multiprocessing.connection import Client, Listener
from typing import Dict, List, Optional

import numpy as np
import torch

from ..cluster import tf_config as tfc
from .strategy import Strategy, _pick_device

log = logging.getLogger("tfde_b200")
_AUTH = b"tfde_b200-ps"


class _ServerState:
    def __init__(self):
        self.vars: Dict[str, np.ndarray] = {}
        self.slots: Dict[str, List[np.ndarray]] = {}
        self.opt = {"kind": "sgd", "momentum": 0.0, "nesterov": False, "beta_1": 0.9, "beta_2": 0.999,
                    "epsilon": 1e-7}
        self.global_step = 0
        self.initialized = False
        self.done = set()
        self.shutdown = threading.Event()
        self.lock = threading.Lock()

    def apply(self, name, g, lr):
        w = self.vars[name]
        kind = self.opt["kind"]
        if kind == "sgd":
            w -= lr * g                      # in place; concurrent workers interleave (async SGD)
        elif kind in ("momentum", "nesterov"):
            v = self.slots.setdefault(name, [np.zeros_like(w)])[0]
            mu = self.opt["momentum"]
            v *= mu
            v -= lr * g
            w += (mu * v - lr * g) if kind == "nesterov" else v
        else:  # adam, bias-corrected with the global step
            sl = self.slots.setdefault(name, [np.zeros_like(w), np.zeros_like(w)])
            b1, b2, eps = self.opt["beta_1"], self.opt["beta_2"], self.opt["epsilon"]
            t = self.global_step + 1
            sl[0] *= b1; sl[0] += (1 - b1) * g
            sl[1] *= b2; sl[1] += (1 - b2) * g * g
            w -= lr * np.sqrt(1 - b2 ** t) / (1 - b1 ** t) * sl[0] / (np.sqrt(sl[1]) + eps)


def _serve_connection(conn, st: _ServerState, counts_steps: bool):
    try:
        while not st.shutdown.is_set():
            try:
                msg = conn.recv()
            except (EOFError, ConnectionResetError, OSError):
                return
            op = msg[0]
            if op == "init":
                _, arrays, opt_cfg, step = msg
                with st.lock:
                    fresh = not st.initialized
                    if fresh:
                        st.vars.update({k: np.array(v, dtype=np.float32, copy=True) for k, v in arrays.items()})
                        if opt_cfg:
                            st.opt.update(opt_cfg)
                        st.global_step = int(step)
                        st.initialized = True
                conn.send(("ok", fresh))
            elif op == "wait_init":
                t0 = time.time()
                while not st.initialized and time.time() - t0 < msg[1] and not st.shutdown.is_set():
                    time.sleep(0.01)
                conn.send(("ok", st.initialized))
            elif op == "pull":
                conn.send(("ok", {k: st.vars[k] for k in msg[1] if k in st.vars}, st.global_step))
            elif op == "push":
                _, grads, lr, assigns = msg
                for k, g in grads.items():
                    st.apply(k, g, lr)
                for k, v in (assigns or {}).items():
                    st.vars[k][...] = v
                if counts_steps:
                    with st.lock:
                        st.global_step += 1
                conn.send(("ok", st.global_step))
            elif op == "step":
                conn.send(("ok", st.global_step))
            elif op == "set_step":
                with st.lock:
                    st.global_step = max(st.global_step, int(msg[1]))
                conn.send(("ok", st.global_step))
            elif op == "done":
                with st.lock:
                    st.done.add(msg[1])
                conn.send(("ok", len(st.done)))
            elif op == "num_done":
                conn.send(("ok", len(st.done)))
            elif op == "shutdown":
                st.shutdown.set()
                conn.send(("ok", True))
                return
            else:
                conn.send(("error", f"unknown op {op!r}"))
    finally:
        try:
            conn.close()
        except Exception:
            pass


class ParameterServer:
    """One ``ps`` task. ``serve_forever`` returns after a ``shutdown`` message (sent by the chief when
    training ends) -- TF's ps tasks never return; a clean exit is friendlier to schedulers."""

    def __init__(self, host: str, port: int, index: int):
        self.host, self.port, self.index = host, port, index
        self.state = _ServerState()
        self._listener = None

    def serve_forever(self):
        self._listener = Listener((self.host if self.host != "127.0.0.1" else "127.0.0.1", self.port),
                                  authkey=_AUTH)
        log.info("ps %d serving on %s:%d", self.index, self.host, self.port)
        threads = []

        def acceptor():
            while not self.state.shutdown.is_set():
                try:
                    conn = self._listener.accept()
                except (OSError, EOFError):
                    return
                except Exception:   # failed handshake from a stray client
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
        except Exception:
            pass
        log.info("ps %d shut down at global_step %d", self.index, self.state.global_step)


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
        self._step = 0
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
        for host, port in self.resolver.ps_endpoints():
            while True:
                try:
                    self._clients.append(Client((host, port), authkey=_AUTH))
                    break
                except (ConnectionRefusedError, OSError):
                    if time.time() > deadline:
                        raise TimeoutError(f"parameter server {host}:{port} unreachable")
                    time.sleep(0.1)

    def _call(self, ps: int, *msg):
        c = self._clients[ps]
        c.send(msg)
        rep = c.recv()
        if rep[0] != "ok":
            raise RuntimeError(f"ps {ps}: {rep}")
        return rep[1:]

    def on_train_begin(self, model):
        if not self.distributed:
            return
        self._connect()
        nps = len(self._clients)
        self._owner = {v.name: i % nps for i, v in enumerate(model.variables)}   # round-robin placement
        model.use_cuda_graph = False    # the step contains host-side pulls/pushes
        opt = model.optimizer
        cfg = {"kind": opt.kind, "momentum": opt.momentum, "nesterov": getattr(opt, "nesterov", False),
               "beta_1": opt.beta_1, "beta_2": opt.beta_2, "epsilon": opt.epsilon}
        if self.is_chief:
            model.sync()
            named = model.named_weights()
            for ps in range(nps):
                mine = {k: v for k, v in named.items() if self._owner[k] == ps}
                self._call(ps, "init", mine, cfg, opt.iterations)
        else:
            for ps in range(nps):
                ok, = self._call(ps, "wait_init", float(os.environ.get("TFDE_RENDEZVOUS_TIMEOUT", "120")))
                if not ok:
                    raise TimeoutError("chief never initialised the parameter servers")
        self.pull(model)

    def global_step(self) -> int:
        if not self.distributed:
            return self._step
        self._connect()
        self._step, = self._call(0, "step")
        return self._step

    def pull(self, model):
        nps = len(self._clients)
        for ps in range(nps):
            names = [v.name for v in model.variables if self._owner[v.name] == ps]
            if not names:
                continue
            arrays, step = self._call(ps, "pull", names)
            if ps == 0:
                self._step = step
            for v in model.variables:
                if v.name in arrays:
                    v.value.copy_(torch.from_numpy(arrays[v.name]).to(v.value.device, non_blocking=True))

    def push(self, model):
        model.sync()
        lr = model.optimizer.learning_rate
        nps = len(self._clients)
        for ps in range(nps):
            grads = {v.name: v.grad.detach().cpu().numpy() for v in model.variables
                     if v.trainable and self._owner[v.name] == ps}
            assigns = {v.name: v.value.detach().cpu().numpy() for v in model.variables
                       if not v.trainable and self._owner[v.name] == ps}
            step, = self._call(ps, "push", grads, lr, assigns)
            if ps == 0:
                self._step = step

    def after_step(self, _local_step) -> int:
        return self._step if self.distributed else _local_step + 1

    def on_train_end(self, model):
        if not self.distributed:
            return
        self._call(0, "done", self.worker_id)
        if self.is_chief:
            # wait (bounded) for the other workers, then release the ps tasks
            n_workers = len(self.resolver.cluster.training_tasks())
            deadline = time.time() + float(os.environ.get("TFDE_PS_DRAIN_TIMEOUT", "30"))
            while time.time() < deadline:
                n, = self._call(0, "num_done")
                if n >= n_workers:
                    break
                time.sleep(0.05)
            self.pull(model)   # final weights for the last checkpoint / export
            for ps in range(len(self._clients)):
                try:
                    self._call(ps, "shutdown")
                except Exception:
                    pass
        for c in self._clients:
            try:
                c.close()
            except Exception:
                pass
        self._clients = []
