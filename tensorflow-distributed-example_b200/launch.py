"""Local launcher: one process per replica, each with its own TF_CONFIG.

TensorFlow's MirroredStrategy drives every local GPU from one process; this framework runs one process
per GPU, so ``MirroredStrategy`` / ``MultiWorkerMirroredStrategy`` jobs on one host are started with::

    python -m tfde_b200.launch --nproc 8 examples/distributed_with_keras.py
    python -m tfde_b200.launch --nproc 2 --ps 1 examples/mnist_keras_distributed.py --working-dir /tmp/w

which is the scripted form of "export a different TF_CONFIG in N shells" (SURVEY.md section 4). Every child
gets ``TF_CONFIG`` (cluster of localhost:port tasks + its own task), ``LOCAL_RANK`` and -- for the
reference's YARN-style scripts -- ``CLUSTER_SPEC`` / ``JOB_NAME`` / ``TASK_INDEX``. torchrun works too
(the strategies also understand RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT).
"""
from __future__ import annotations

import argparse
import json
import os
import socket
import subprocess
import sys
import time


def free_ports(n: int):
    socks, ports = [], []
    for _ in range(n):
        s = socket.socket()
        s.bind(("127.0.0.1", 0))
        socks.append(s)
        ports.append(s.getsockname()[1])
    for s in socks:
        s.close()
    return ports


def build_cluster(nproc: int, n_ps: int = 0, chief: bool = False, evaluator: bool = False):
    ports = free_ports(nproc + n_ps + (1 if evaluator else 0))
    it = iter(ports)
    cluster, tasks = {}, []
    n_workers = nproc - (1 if chief else 0)
    if chief:
        cluster["chief"] = [f"127.0.0.1:{next(it)}"]
        tasks.append(("chief", 0))
    if n_workers:
        cluster["worker"] = [f"127.0.0.1:{next(it)}" for _ in range(n_workers)]
        tasks += [("worker", i) for i in range(n_workers)]
    if n_ps:
        cluster["ps"] = [f"127.0.0.1:{next(it)}" for _ in range(n_ps)]
        tasks += [("ps", i) for i in range(n_ps)]
    if evaluator:
        cluster["evaluator"] = [f"127.0.0.1:{next(it)}"]
        tasks.append(("evaluator", 0))
    return cluster, tasks


def launch(script_argv, nproc=2, n_ps=0, chief=False, evaluator=False, env_style="tf_config",
           extra_env=None, timeout=None, capture=False):
    """Starts the processes, waits for them, returns the list of (task, returncode[, stdout])."""
    cluster, tasks = build_cluster(nproc, n_ps, chief, evaluator)
    procs = []
    local_rank = 0
    for job, idx in tasks:
        env = dict(os.environ)
        env.update(extra_env or {})
        if env_style == "cluster_spec":
            env["CLUSTER_SPEC"] = json.dumps(cluster)
            env["JOB_NAME"], env["TASK_INDEX"] = job, str(idx)
            env.pop("TF_CONFIG", None)
        else:
            env["TF_CONFIG"] = json.dumps({"cluster": cluster, "task": {"type": job, "index": idx}})
        for k in ("RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT"):
            env.pop(k, None)
        if job in ("chief", "master", "worker"):
            env["LOCAL_RANK"] = str(local_rank)
            local_rank += 1
        else:
            env["LOCAL_RANK"] = "0"
        kw = dict(stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) if capture else {}
        procs.append(((job, idx), subprocess.Popen([sys.executable] + list(script_argv), env=env, **kw)))
    results = []
    deadline = None if timeout is None else time.time() + timeout
    for task, p in procs:
        try:
            out, _ = p.communicate(timeout=None if deadline is None else max(1.0, deadline - time.time()))
        except subprocess.TimeoutExpired:
            p.kill()
            out, _ = p.communicate()
        results.append((task, p.returncode, out) if capture else (task, p.returncode))
    return results


def main():
    ap = argparse.ArgumentParser(prog="python -m tfde_b200.launch")
    ap.add_argument("--nproc", type=int, default=2, help="training processes (chief + workers)")
    ap.add_argument("--ps", type=int, default=0, help="parameter-server tasks")
    ap.add_argument("--chief", action="store_true", help="name the first training task 'chief'")
    ap.add_argument("--evaluator", action="store_true", help="add an evaluator task")
    ap.add_argument("--env-style", choices=["tf_config", "cluster_spec"], default="tf_config")
    ap.add_argument("script", nargs=argparse.REMAINDER)
    a = ap.parse_args()
    if not a.script:
        ap.error("missing script")
    res = launch(a.script, a.nproc, a.ps, a.chief, a.evaluator, a.env_style)
    bad = [r for r in res if r[1] != 0]
    for task, rc in res:
        print(f"[launch] {task[0]}:{task[1]} exited with {rc}")
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
