"""Cluster discovery: TF_CONFIG and the CLUSTER_SPEC / JOB_NAME / TASK_INDEX shim.

Reference behaviour being reproduced:

* ``TF_CONFIG`` is a JSON object ``{"cluster": {job: [host:port, ...]}, "task": {"type", "index"}}``
  read by the strategy constructors (distributed_with_keras.py:16 -- implicitly, inside TensorFlow).
* ``CLUSTER_SPEC`` (JSON), ``JOB_NAME`` and ``TASK_INDEX`` are turned into ``TF_CONFIG`` and printed
  (mnist_keras_distributed.py:221-233).
* per-role "device filters": master sees ps+master, worker i sees ps + itself
  (mnist_keras_distributed.py:165-189).

B200-native mapping: one process per GPU. Roles that *train synchronously* (chief/master + worker)
are numbered into ranks ``0..world-1`` in TF's order (chief first, then workers); the rendezvous
endpoint of the control plane (a ``torch.distributed`` TCP store -- no gradient byte goes through it)
is rank 0's ``host:port`` from the cluster spec. ``ps`` and ``evaluator`` tasks are not part of the
collective group.
"""
from __future__ import annotations

import json
import os
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

TRAIN_JOBS = ("chief", "master", "worker")
ALL_JOBS = ("chief", "master", "worker", "ps", "evaluator")


class ClusterConfigError(ValueError):
    pass


@dataclass
class ClusterSpec:
    jobs: Dict[str, List[str]] = field(default_factory=dict)

    def num_tasks(self, job: str) -> int:
        return len(self.jobs.get(job, []))

    def task_address(self, job: str, index: int) -> str:
        try:
            return self.jobs[job][index]
        except (KeyError, IndexError):
            raise ClusterConfigError(f"no task {job}:{index} in cluster {self.jobs}")

    def as_dict(self):
        return {k: list(v) for k, v in self.jobs.items()}

    def training_tasks(self) -> List[Tuple[str, int]]:
        """(job, index) of every task that takes part in synchronous training, in rank order."""
        out = []
        for job in TRAIN_JOBS:
            out += [(job, i) for i in range(self.num_tasks(job))]
        return out


@dataclass
class TaskInfo:
    type: str = "worker"
    index: int = 0


@dataclass
class ClusterResolver:
    """Parsed view of TF_CONFIG (the analogue of tf.distribute.cluster_resolver.TFConfigClusterResolver)."""
    cluster: ClusterSpec
    task: TaskInfo
    raw: Optional[dict] = None

    # ---- roles -------------------------------------------------------------------------------
    @property
    def distributed(self) -> bool:
        return bool(self.cluster.jobs)

    @property
    def is_chief(self) -> bool:
        """TF rule: the 'chief' (or legacy 'master') task if the job exists, else worker 0."""
        if not self.distributed:
            return True
        if self.cluster.num_tasks("chief") or self.cluster.num_tasks("master"):
            return self.task.type in ("chief", "master")
        return self.task.type == "worker" and self.task.index == 0

    @property
    def is_ps(self) -> bool:
        return self.task.type == "ps"

    @property
    def is_evaluator(self) -> bool:
        return self.task.type == "evaluator"

    # ---- ranks -------------------------------------------------------------------------------
    @property
    def world_size(self) -> int:
        return max(1, len(self.cluster.training_tasks())) if self.distributed else 1

    @property
    def rank(self) -> int:
        if not self.distributed:
            return 0
        tasks = self.cluster.training_tasks()
        try:
            return tasks.index((self.task.type, self.task.index))
        except ValueError:
            return -1  # ps / evaluator: not a member of the synchronous group

    @property
    def master_endpoint(self) -> Tuple[str, int]:
        tasks = self.cluster.training_tasks()
        if not tasks:
            return "127.0.0.1", 29500
        host, port = split_hostport(self.cluster.task_address(*tasks[0]))
        return host, port

    def ps_endpoints(self) -> List[Tuple[str, int]]:
        return [split_hostport(a) for a in self.cluster.jobs.get("ps", [])]

    def num_workers(self) -> int:
        return self.world_size

    # ---- device filters (C5) -----------------------------------------------------------------
    def device_filters(self) -> Optional[List[str]]:
        return device_filters_from_tf_config(self.raw or {})


def split_hostport(addr: str) -> Tuple[str, int]:
    if ":" not in addr:
        raise ClusterConfigError(f"task address {addr!r} is not host:port")
    host, port = addr.rsplit(":", 1)
    if host in ("localhost", ""):
        host = "127.0.0.1"
    return host, int(port)


def parse_tf_config(text: Optional[str]) -> ClusterResolver:
    """Parses a TF_CONFIG JSON string (``None``/empty -> a single local worker)."""
    if not text:
        return ClusterResolver(ClusterSpec({}), TaskInfo("worker", 0), raw={})
    try:
        cfg = json.loads(text)
    except json.JSONDecodeError as e:
        raise ClusterConfigError(f"TF_CONFIG is not valid JSON: {e}") from e
    if not isinstance(cfg, dict):
        raise ClusterConfigError("TF_CONFIG must be a JSON object")
    cluster = cfg.get("cluster", {}) or {}
    for job, tasks in cluster.items():
        if job not in ALL_JOBS:
            raise ClusterConfigError(f"unknown job {job!r} in TF_CONFIG cluster (expected one of {ALL_JOBS})")
        if not isinstance(tasks, list) or not all(isinstance(t, str) for t in tasks):
            raise ClusterConfigError(f"job {job!r} must map to a list of host:port strings")
    task = cfg.get("task", {}) or {}
    ttype = task.get("type", "worker")
    tindex = int(task.get("index", 0))
    spec = ClusterSpec({k: list(v) for k, v in cluster.items()})
    if cluster:
        if ttype not in cluster and ttype != "evaluator":
            raise ClusterConfigError(f"task type {ttype!r} not present in cluster {list(cluster)}")
        if ttype in cluster and not (0 <= tindex < len(cluster[ttype])):
            raise ClusterConfigError(f"task index {tindex} out of range for job {ttype!r}")
    return ClusterResolver(spec, TaskInfo(ttype, tindex), raw=cfg)


def resolve_from_env(environ=None) -> ClusterResolver:
    env = os.environ if environ is None else environ
    return parse_tf_config(env.get("TF_CONFIG"))


def synthesize_tf_config_from_cluster_spec(environ=None, verbose=True) -> Optional[str]:
    """CLUSTER_SPEC/JOB_NAME/TASK_INDEX -> TF_CONFIG (mnist_keras_distributed.py:221-233).

    Sets ``environ['TF_CONFIG']`` and prints the same two status lines as the reference. Unlike the
    reference this does not leave ``job_type`` unbound when CLUSTER_SPEC is absent.
    """
    env = os.environ if environ is None else environ
    spec = env.get("CLUSTER_SPEC")
    if not spec:
        if verbose:
            print("Distribution is not enabled")
        return None
    try:
        cluster = json.loads(spec)
    except json.JSONDecodeError as e:
        raise ClusterConfigError(f"CLUSTER_SPEC is not valid JSON: {e}") from e
    if "TASK_INDEX" not in env or "JOB_NAME" not in env:
        raise ClusterConfigError("CLUSTER_SPEC is set but TASK_INDEX / JOB_NAME are missing")
    tf_config = json.dumps({"cluster": cluster,
                            "task": {"type": env["JOB_NAME"], "index": int(env["TASK_INDEX"])}})
    env["TF_CONFIG"] = tf_config
    if verbose:
        print("Distribution enabled: ", tf_config)
    return tf_config


def device_filters_from_tf_config(cfg: dict) -> Optional[List[str]]:
    """master -> ['/job:ps', '/job:master']; worker i -> ['/job:ps', '/job:worker/task:i']; else None
    (mnist_keras_distributed.py:176-189). In this framework the filter list decides which peers a
    process maps memory from / opens sockets to in parameter-server mode."""
    task = cfg.get("task") if cfg else None
    if not task or "type" not in task or "index" not in task:
        return None
    if task["type"] == "master":
        return ["/job:ps", "/job:master"]
    if task["type"] == "worker":
        return ["/job:ps", "/job:worker/task:%d" % int(task["index"])]
    return None


def make_tf_config(cluster: Dict[str, List[str]], task_type: str, task_index: int) -> str:
    return json.dumps({"cluster": cluster, "task": {"type": task_type, "index": task_index}})


def tf_config_from_torchrun_env(environ=None) -> Optional[str]:
    """When launched by torchrun (RANK/WORLD_SIZE/MASTER_ADDR/MASTER_PORT) synthesize the equivalent
    TF_CONFIG so the rest of the stack has one source of truth."""
    env = os.environ if environ is None else environ
    if "RANK" not in env or "WORLD_SIZE" not in env:
        return None
    world = int(env["WORLD_SIZE"])
    addr = env.get("MASTER_ADDR", "127.0.0.1")
    port = int(env.get("MASTER_PORT", "29500"))
    workers = [f"{addr}:{port + i}" for i in range(world)]
    return make_tf_config({"worker": workers}, "worker", int(env["RANK"]))
