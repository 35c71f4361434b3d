"""Distribution strategies: the ``tf.distribute`` surface of the reference, B200-first.

* :class:`MultiWorkerMirroredStrategy` -- synchronous data parallel across processes; no-arg ctor that
  reads ``TF_CONFIG`` (distributed_with_keras.py:16), ``scope()`` around dataset/model construction
  (distributed_with_keras.py:51-58).
* :class:`MirroredStrategy` -- synchronous data parallel inside one host
  (mnist_keras_distributed.py:243, used there as ``eval_distribute``).
* :class:`ParameterServerStrategy` -- asynchronous data parallel with ps/worker/master roles
  (mnist_keras_distributed.py:241-242, tf2_mnist_distributed.py:189); see :mod:`.ps`.

Execution model (differs from TF on purpose): ONE PROCESS PER GPU. ``num_replicas_in_sync`` is the
number of processes in the synchronous group. ``torch.distributed`` (gloo on CPU, NCCL or gloo on
GPU) is the *control plane only* -- rendezvous, handle exchange, barriers. On GPUs the per-step
gradient traffic is the hand-written peer-to-peer kernel in ``csrc/kernels/allreduce_p2p.cu`` fused with
the optimizer (:mod:`.comm`); on CPU it is a gloo all-reduce (plumbing/test path, BASELINE.json
config 1).

Semantics reproduced: global-batch datasets are split evenly per replica;
loss = sum(per-example)/global_batch so SUM-reduced gradients equal the single-device gradient;
initial variables are broadcast from the chief; replicas stay bit-identical; checkpoints, summaries
and TensorBoard are chief-only.
"""
from __future__ import annotations

import contextlib
import datetime
import logging
import os
import threading
from typing import Optional

import torch
import torch.distributed as dist

from ..cluster import tf_config as tfc

log = logging.getLogger("tfde_b200")

_state = threading.local()


def get_strategy() -> "Strategy":
    s = getattr(_state, "stack", None)
    if s:
        return s[-1]
    return _default_strategy()


_DEFAULT = None


def _default_strategy():
    global _DEFAULT
    if _DEFAULT is None:
        _DEFAULT = Strategy()
    return _DEFAULT


def has_strategy() -> bool:
    return bool(getattr(_state, "stack", None))


def _pick_device(local_rank: int, device: Optional[str]) -> torch.device:
    forced = device if device is not None else os.environ.get("TFDE_DEVICE")
    if forced:
        d = torch.device(forced)
        if d.type == "cuda" and d.index is None:
            d = torch.device("cuda", local_rank % max(torch.cuda.device_count(), 1))
        return d
    if torch.cuda.is_available():
        return torch.device("cuda", local_rank % max(torch.cuda.device_count(), 1))
    return torch.device("cpu")


class Strategy:
    """Default (single replica) strategy and base class."""

    def __init__(self, device: Optional[str] = None):
        self.rank = 0
        self.world_size = 1
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        self.device = _pick_device(self.local_rank, device)
        if self.device.type == "cuda":
            torch.cuda.set_device(self.device)
        self.resolver = tfc.parse_tf_config(None)
        self.group = None
        self._comm = None

    # ---- tf.distribute.Strategy surface ------------------------------------------------------
    @property
    def num_replicas_in_sync(self) -> int:
        return self.world_size

    @property
    def replica_id(self) -> int:
        return self.rank

    @property
    def is_chief(self) -> bool:
        return self.rank == 0

    @property
    def synchronous(self) -> bool:
        return True

    @contextlib.contextmanager
    def scope(self):
        stack = getattr(_state, "stack", None)
        if stack is None:
            stack = _state.stack = []
        stack.append(self)
        try:
            yield self
        finally:
            stack.pop()

    def experimental_distribute_dataset(self, dataset):
        from ..data.dataset import DistributedDataset
        return DistributedDataset(dataset, self)

    def barrier(self):
        if self.world_size > 1:
            dist.barrier(group=self.group)

    # ---- collectives used by the training engine -----------------------------------------------
    def comm(self):
        """Gradient communication backend, created lazily once the model's arenas exist."""
        if self._comm is None:
            from .comm import make_comm
            self._comm = make_comm(self)
        return self._comm

    def broadcast_(self, tensor: torch.Tensor, src: int = 0):
        """Initial-value broadcast from the chief (N3 / G13)."""
        if self.world_size > 1:
            self.comm().broadcast_(tensor, src)
        return tensor

    def all_reduce_sum_(self, tensor: torch.Tensor):
        """Small-tensor SUM (metric aggregation, G14)."""
        if self.world_size > 1:
            self.comm().all_reduce_sum_(tensor)
        return tensor

    def shutdown(self):
        if self._comm is not None:
            self._comm.close()
            self._comm = None

    def __repr__(self):
        return (f"{type(self).__name__}(rank={self.rank}, world={self.world_size}, "
                f"device={self.device}, chief={self.is_chief})")


def _init_process_group(rank: int, world: int, host: str, port: int, device: torch.device):
    """Control-plane rendezvous. Idempotent: reuses an already initialised default group (torchrun)."""
    if device.type == "cpu" and "OMP_NUM_THREADS" not in os.environ:
        # several CPU workers on one host: split the cores instead of oversubscribing them N-fold
        torch.set_num_threads(max(1, (os.cpu_count() or 1) // max(world, 1)))
    if dist.is_initialized():
        return
    backend = os.environ.get("TFDE_CONTROL_BACKEND")
    if backend is None:
        backend = "gloo" if device.type == "cpu" else "cpu:gloo,cuda:nccl"
    timeout = datetime.timedelta(seconds=int(os.environ.get("TFDE_RENDEZVOUS_TIMEOUT", "300")))
    init = f"tcp://{host}:{port}"
    log.info("rendezvous %s rank %d/%d backend %s", init, rank, world, backend)
    kwargs = {}
    if device.type == "cuda":
        kwargs["device_id"] = device
    dist.init_process_group(backend=backend, init_method=init, rank=rank, world_size=world,
                            timeout=timeout, **kwargs)


class MultiWorkerMirroredStrategy(Strategy):
    """Synchronous data parallel over the processes named by TF_CONFIG (or torchrun's env).

    ``MultiWorkerMirroredStrategy()`` mirrors the reference's no-arg construction. Every variable
    created under ``scope()`` is broadcast from the chief, every step SUM-reduces the gradients and
    applies the identical update on every replica.
    """

    def __init__(self, cluster_resolver: Optional[tfc.ClusterResolver] = None, device: Optional[str] = None):
        if cluster_resolver is None:
            text = os.environ.get("TF_CONFIG") or tfc.tf_config_from_torchrun_env()
            cluster_resolver = tfc.parse_tf_config(text)
        self.resolver = cluster_resolver
        r = cluster_resolver
        if r.distributed and r.rank < 0:
            raise tfc.ClusterConfigError(
                f"task {r.task.type}:{r.task.index} is not a training task; "
                "MultiWorkerMirroredStrategy runs on chief/master/worker tasks only")
        self.rank = r.rank if r.distributed else 0
        self.world_size = r.world_size
        self.local_rank = int(os.environ.get("LOCAL_RANK", str(self.rank)))
        self.device = _pick_device(self.local_rank, device)
        if self.device.type == "cuda":
            torch.cuda.set_device(self.device)
        self.group = None
        self._comm = None
        if self.world_size > 1:
            if "MASTER_ADDR" in os.environ and "MASTER_PORT" in os.environ and "RANK" in os.environ:
                host, port = os.environ["MASTER_ADDR"], int(os.environ["MASTER_PORT"])
            else:
                host, port = r.master_endpoint
            _init_process_group(self.rank, self.world_size, host, port, self.device)

    @property
    def is_chief(self) -> bool:
        return self.resolver.is_chief if self.resolver.distributed else True


class MirroredStrategy(MultiWorkerMirroredStrategy):
    """Synchronous data parallel across the GPUs of one host (mnist_keras_distributed.py:243 uses it as
    ``eval_distribute``).

    TF drives all local GPUs from one process; here TRAINING runs one process per GPU: start them with
    ``python -m tfde_b200.launch --nproc N script.py`` or torchrun, and inside such a group this class behaves
    exactly like :class:`MultiWorkerMirroredStrategy`. Everywhere else -- started stand-alone, or constructed inside
    a task of a parameter-server / Estimator cluster (``TF_CONFIG`` with ``ps`` or ``evaluator`` jobs, which is
    where the reference builds it) -- it is LOCAL: it never joins or waits for the cluster's rendezvous, trains as a
    single replica, and EVALUATES across the local GPUs from this one process (``local_devices``; per-device model
    replicas, each batch split between them, (loss_sum, correct, count) summed -- ``Sequential.evaluate``).

    ``devices``: e.g. ``["/gpu:0", "/gpu:1"]`` / ``["cuda:0", "cuda:1"]``; default: every visible GPU
    (``TFDE_MIRRORED_DEVICES=0,1`` overrides)."""

    def __init__(self, devices=None, device: Optional[str] = None):
        text = tfc.tf_config_from_torchrun_env()
        if text is None:
            env = os.environ.get("TF_CONFIG")
            if env:
                r = tfc.parse_tf_config(env)
                sync_group = (r.distributed and r.rank >= 0 and not r.cluster.num_tasks("ps")
                              and not r.cluster.num_tasks("evaluator"))
                if sync_group:
                    text = env            # the launcher's synthesized cluster of local training tasks
        super().__init__(tfc.parse_tf_config(text), device=device)
        self.devices = devices

    @property
    def local_devices(self):
        """Devices one process evaluates on. A member of a multi-process group owns exactly its own GPU."""
        if self.world_size > 1 or self.device.type != "cuda":
            return [self.device]
        spec = self.devices
        if spec is None and os.environ.get("TFDE_MIRRORED_DEVICES"):
            spec = [f"cuda:{int(i)}" for i in os.environ["TFDE_MIRRORED_DEVICES"].split(",") if i.strip() != ""]
        if spec is None:
            n = torch.cuda.device_count()
            rest = [torch.device("cuda", i) for i in range(n) if i != self.device.index]
            return [self.device] + rest
        out = []
        for d in spec:
            name = str(d).lower().replace("/device:", "").replace("/", "").replace("gpu", "cuda")
            out.append(torch.device(name))
        return out or [self.device]


class OneDeviceStrategy(Strategy):
    pass
