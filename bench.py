#!/usr/bin/env python
"""Headline benchmark: MNIST CNN training throughput (images/sec) at N B200s, weak scaling.

Metric and config follow BASELINE.json: the Keras MNIST CNN of the reference
(distributed_with_keras.py:33-43 == the model_fn of tf2_mnist_distributed.py:66-72: Conv32-Pool-Dense64-
Dense10, SCCE from logits, SGD lr 0.001), per-replica batch 64 (distributed_with_keras.py:13), global batch
64 x N under the MirroredStrategy/MultiWorkerMirroredStrategy equivalent, fp32 like the reference
(distributed_with_keras.py:21), synthetic 28x28x1 data, random-init weights.

    python bench.py --gpus 1 --steps 200 --warmup 20
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29511 bench.py --gpus 8 --steps 200 --warmup 20

Two numbers per run, printed as ONE JSON line by rank 0:

* ``value`` -- device-timed: K steps back to back between one pair of CUDA events on the model's stream, launched
  through the public API (``model.train_on_batches``) from device-resident inputs; every step reads a distinct
  batch from a pool larger than L2 (evicted before each repeat); repeated ``--repeats`` times, median of the
  per-repeat max over ranks. A step = forward, loss, backward, cross-replica gradient reduction, SGD update and
  metric accumulation (for this model: one cooperative kernel, ``csrc/kernels/fused_step.cu``).
* ``e2e`` -- wall clock through the same public API from HOST batches: every step's uint8 images + int64 labels
  are copied host->device from pinned memory (one cudaMemcpyAsync each per step) and every step's loss is read back
  on the host, all inside the timed region.

``--impl reference`` must run the unmodified TensorFlow reference; TensorFlow cannot be installed on
this image (no wheel, no network), so that arm reports ``unavailable``. ``--impl torch`` runs the
stand-in baseline of BASELINE.md section 4 (PyTorch eager + cuDNN/cuBLAS + DDP/NCCL).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "mnist_cnn_train_images_per_sec"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "torch"])
    ap.add_argument("--model", default="a", choices=["a", "b"])
    ap.add_argument("--batch-per-gpu", type=int, default=64)
    ap.add_argument("--optimizer", default="sgd", choices=["sgd", "momentum", "adam"])
    ap.add_argument("--dtype", default="fp32", choices=["fp32", "bf16"],
                    help="bf16 = mixed_bfloat16 policy: large Dense layers on the tcgen05 GEMM (BASELINE.json config 3)")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-flush", action="store_true")
    ap.add_argument("--repeats", type=int, default=5, help="timed regions of K steps each; the median is reported")
    ap.add_argument("--steps-per-execution", type=int, default=0,
                    help="steps handed to the device per launch (Keras steps_per_execution); 0 = min(64, steps) with streamed inputs (Model A), else min(16, steps/4) so that the copies of one execution overlap the kernel of the previous one")
    ap.add_argument("--no-check", action="store_true", help="skip the replica / CPU-oracle correctness check")
    ap.add_argument("--l2", default="inputs", choices=["inputs", "flush"],
                    help="timing protocol for `value`: distinct inputs from a pool larger than L2 (default) or "
                         "an L2 flush before every step")
    return ap.parse_args()


class ClockSampler:
    """SM clock + throttle-reason sampling DURING the timed region (B200_PROFILING.md clocks line).

    Reads the counters nvidia-smi prints (clocks.sm, clocks.max.sm, clocks_event_reasons.*) through NVML
    in-process, every few milliseconds, from a thread. Spawning ``nvidia-smi -lms`` next to a 40 ms timed
    region does not work: its start-up (NVML attach to every GPU) is longer than the region and perturbs
    kernel launches on all ranks, so the samples land after the fact. ``begin()``/``end()`` mark the timed
    window; only samples inside it are reported. Falls back to an nvidia-smi subprocess without pynvml."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown",
               0x4: "sw_power_cap", 0x80: "hw_power_brake_slowdown"}

    def __init__(self, gpu_index=0, period_s=0.0005):
        self.gpu_index, self.period_s = gpu_index, period_s
        self.proc = None
        self.lines = []
        self.samples = []          # (t, sm_mhz, reason_mask)
        self.windows = []
        self._stop = threading.Event()
        self._nvml = None
        self.max_mhz = None

    def _nvml_handle(self):
        import pynvml
        pynvml.nvmlInit()
        try:
            import torch
            uuid = str(torch.cuda.get_device_properties(self.gpu_index).uuid)
            h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
        except Exception:
            h = pynvml.nvmlDeviceGetHandleByIndex(self.gpu_index)
        return pynvml, h

    def start(self):
        try:
            self._nvml, self._h = self._nvml_handle()
            self.max_mhz = float(self._nvml.nvmlDeviceGetMaxClockInfo(self._h, self._nvml.NVML_CLOCK_SM))
            self._t = threading.Thread(target=self._poll, daemon=True)
            self._t.start()
            return
        except Exception:
            self._nvml = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "20",
                 "-i", str(self.gpu_index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self._t = threading.Thread(target=self._read, daemon=True)
            self._t.start()
        except Exception:
            self.proc = None

    def begin(self):
        self._w0 = time.perf_counter()

    def end(self):
        self.windows.append((self._w0, time.perf_counter()))

    def _poll(self):
        n = self._nvml
        reasons = getattr(n, "nvmlDeviceGetCurrentClocksEventReasons", None) or n.nvmlDeviceGetCurrentClocksThrottleReasons
        while not self._stop.is_set():
            try:
                self.samples.append((time.perf_counter(), float(n.nvmlDeviceGetClockInfo(self._h, n.NVML_CLOCK_SM)),
                                     int(reasons(self._h))))
            except Exception:
                pass
            self._stop.wait(self.period_s)

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self._nvml is not None:
            self._stop.set()
            self._t.join(timeout=1)
            inside = [x for x in self.samples if any(a <= x[0] <= b for a, b in self.windows)] if self.windows else []
            use = inside or self.samples
            mask = 0
            for x in use:
                mask |= x[2]
            return {"sm_mhz": statistics.median([x[1] for x in use]) if use else None, "sm_max_mhz": self.max_mhz,
                    "samples": len(use), "samples_inside_timed_region": len(inside), "source": "nvml",
                    "reasons": sorted(v for k, v in self.REASONS.items() if mask & k)}
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.05)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [s.strip() for s in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "source": "nvidia-smi", "reasons": sorted(reasons)}


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def max_over_ranks(val: float, world: int) -> float:
    if world == 1:
        return val
    import torch
    import torch.distributed as dist
    t = torch.tensor([val], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def barrier(world):
    if world > 1:
        import torch.distributed as dist
        dist.barrier()


def make_batches(n_batches, batch, seed):
    """Synthetic MNIST-shaped pinned host batches: uint8 [B,28,28,1] images, int64 [B] labels."""
    import torch
    g = torch.Generator().manual_seed(seed)
    xs = [torch.randint(0, 256, (batch, 28, 28, 1), generator=g, dtype=torch.uint8).pin_memory()
          for _ in range(n_batches)]
    ys = [torch.randint(0, 10, (batch,), generator=g, dtype=torch.int64).pin_memory() for _ in range(n_batches)]
    return xs, ys


def _weights_digest(model):
    import hashlib
    h = hashlib.sha256()
    for w in model.get_weights():
        h.update(w.tobytes())
    return h.hexdigest()


def correctness_check(tf, strategy, args, rank, world, steps=3):
    """What the driver's records should carry besides a rate: after `steps` optimizer steps of a fresh model on
    all replicas, (a) every replica holds bit-identical weights, (b) they match, to fp32 tolerance, ONE process
    training on the concatenated global batch through the PyTorch CPU restatement (ops/reference.py)."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from tfde_b200.parallel.strategy import Strategy
    B = args.batch_per_gpu
    gb = B * world

    def build(opt_seed=77):
        opt = {"sgd": lambda: tf.keras.optimizers.SGD(0.05),
               "momentum": lambda: tf.keras.optimizers.SGD(0.05, momentum=0.9),
               "adam": lambda: tf.keras.optimizers.Adam(0.002)}[args.optimizer]()
        if args.model == "a":
            return tf.models.build_and_compile_cnn_model(optimizer=opt, seed=opt_seed)
        return tf.models.build_and_compile_model_b(optimizer=opt, seed=opt_seed)

    def batches(r):
        g = torch.Generator().manual_seed(900 + r)
        x = torch.randint(0, 256, (steps, B, 28, 28, 1), generator=g, dtype=torch.uint8)
        y = torch.randint(0, 10, (steps, B), generator=g, dtype=torch.int64)
        if args.model == "b":
            x = x.reshape(steps, B, 784).float() / 255.0
        return x, y

    with strategy.scope():
        m = build()
    x, y = batches(rank)
    for k in range(steps):
        m.train_on_batch(x[k], y[k], global_batch=gb)
    m.sync()
    digest = _weights_digest(m)
    identical = True
    if world > 1:
        ds = [None] * world
        dist.all_gather_object(ds, digest)
        identical = all(d == ds[0] for d in ds)
    err = None
    if rank == 0:
        with Strategy(device="cpu").scope():
            mc = build()
        all_xy = [batches(r) for r in range(world)]
        for k in range(steps):
            mc.train_on_batch(torch.cat([a[0][k] for a in all_xy], 0), torch.cat([a[1][k] for a in all_xy], 0),
                              global_batch=gb)
        err = 0.0
        for a, b in zip(m.get_weights(), mc.get_weights()):
            err = max(err, float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-12)))
    return {"replicas_identical": bool(identical), "max_rel_err_vs_single_process": err, "steps": steps,
            "oracle": "PyTorch CPU fp32 restatement, one process, global batch"}


def run_ours(args):
    import torch
    import tfde_b200 as tf
    from tfde_b200.ops import native

    rank, world, local = dist_env()
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a GPU (the CPU path is a plumbing path, not a benchmark)")
    native.load()
    strategy = tf.distribute.MirroredStrategy() if world > 1 else tf.distribute.get_strategy()
    if world > 1 and torch.distributed.is_initialized() is False:
        raise SystemExit("strategy did not initialise the process group")
    B = args.batch_per_gpu
    gb = B * world
    K = args.steps
    W = max(args.warmup, 3)
    R = max(1, args.repeats)
    # steps_per_execution (Keras): the device-timed region feeds device-resident batches, where one launch can take all
    # K steps; the end-to-end region wants >= 4 executions in flight so that the H2D copies of one execution overlap the
    # kernel of the previous one
    spe_dev = args.steps_per_execution if args.steps_per_execution > 0 else max(1, min(64, K))
    from tfde_b200.ops import fused_step as _fs
    if args.steps_per_execution > 0:
        spe = args.steps_per_execution
    elif args.model == "a" and _fs.enabled() and _fs.streamed_inputs():
        # streamed inputs: the whole-step kernel is launched without waiting for the copies and takes up every step when
        # that step's batch has landed, so ONE execution covers the region with the copies hidden under compute
        spe = spe_dev
    else:
        spe = max(1, min(16, K // 4))
    if args.dtype == "bf16":
        tf.keras.mixed_precision.set_global_policy("mixed_bfloat16")
    check = correctness_check(tf, strategy, args, rank, world) if not args.no_check else None
    with strategy.scope():
        opt = {"sgd": lambda: tf.keras.optimizers.SGD(0.001),
               "momentum": lambda: tf.keras.optimizers.SGD(0.001, momentum=0.9),
               "adam": lambda: tf.keras.optimizers.Adam(0.001)}[args.optimizer]()
        if args.model == "a":
            model = tf.models.build_and_compile_cnn_model(optimizer=opt, seed=1234)
            in_shape = (28, 28, 1)
        else:
            model = tf.models.build_and_compile_model_b(optimizer=opt, seed=1234)
            in_shape = (784,)
    model.use_cuda_graph = not args.no_graph
    dev = model.device

    # host data: 8 distinct pinned super-batches of `spe` steps each (uint8 images, int64 labels)
    def host_batches(n, seed):
        g = torch.Generator().manual_seed(seed)
        xs = [torch.randint(0, 256, (n * B, 28, 28, 1), generator=g, dtype=torch.uint8) for _ in range(8)]
        ys = [torch.randint(0, 10, (n * B,), generator=g, dtype=torch.int64).pin_memory() for _ in range(8)]
        if args.model == "b":
            xs = [x.reshape(n * B, 784).float().div_(255.0) for x in xs]
        return [x.pin_memory() for x in xs], ys

    xs, ys = host_batches(spe, 100 + rank)
    tail_n = K % spe
    xt, yt = host_batches(tail_n, 300 + rank) if tail_n else (None, None)

    sampler = ClockSampler(local)
    sampler.start()        # started before the warm-up so that its own start-up is outside the timed regions
    # ---- warm-up (>= 3 steps): first launches / graph capture, every execution shape used below -------------
    for i in range(W):
        model.train_on_batch(xs[i % 8][:B], ys[i % 8][:B], global_batch=gb)
    model.train_on_batches(xs[0], ys[0], steps=spe, global_batch=gb)
    model.train_on_batches(xs[1], ys[1], steps=spe, global_batch=gb)
    if tail_n:
        model.train_on_batches(xt[0], yt[0], steps=tail_n, global_batch=gb)
    model.sync()
    fused = model.fused_plan() is not None

    # ---- device-timed region ---------------------------------------------------------------------------------
    # K steps back to back between one pair of CUDA events, launched through the public API from DEVICE-resident
    # inputs; every step consumes a distinct batch out of a pool larger than L2 (>= 136 MiB, evicted by a 256 MiB
    # write before timing), so inputs come from HBM while weights stay as resident as they are in real training.
    # Repeated R times on different pool ranges; `value` is the median, max over ranks per repeat.
    per = B * 784 * xs[0].element_size() + B * 8
    P = max(K * R, -(-(136 << 20) // per))
    gen = torch.Generator(device=dev).manual_seed(4321 + rank)
    if args.model == "a":
        pool_x = torch.randint(0, 256, (P * B, 28, 28, 1), generator=gen, dtype=torch.uint8, device=dev)
    else:
        pool_x = torch.rand((P * B, 784), generator=gen, dtype=torch.float32, device=dev)
    pool_y = torch.randint(0, 10, (P * B,), generator=gen, dtype=torch.int64, device=dev)
    chunks = [(i, min(spe, K - i)) for i in range(0, K, spe)]
    chunks_dev = [(i, min(spe_dev, K - i)) for i in range(0, K, spe_dev)]

    def timed_device(rep):
        torch.empty(256 << 20, dtype=torch.uint8, device=dev).fill_(rep + 1)      # evict L2
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        base = (rep * K) % max(P - K, 1)
        barrier(world)
        torch.cuda.synchronize()
        c0 = native.launch_count()
        sampler.begin()
        model.record_event(e0)            # both events sit on the stream the model launches its steps on
        for i, n in chunks_dev:
            lo = (base + i) * B
            model.train_on_batches(pool_x[lo:lo + n * B], pool_y[lo:lo + n * B], steps=n, global_batch=gb)
        model.record_event(e1)
        torch.cuda.synchronize()
        sampler.end()
        barrier(world)
        return max_over_ranks(e0.elapsed_time(e1), world), native.launch_count() - c0

    timed_device(R)                    # untimed pass: first launch of the device-resident execution shape
    dev_ms, launches = [], 0
    for rep in range(R):
        ms, launches = timed_device(rep)
        dev_ms.append(ms)
    dev_med = statistics.median(dev_ms)

    # ---- end-to-end region (public API; H2D of every step's inputs from pinned memory, D2H of every step's loss) --
    e2e_s, loss = [], 0.0
    for rep in range(R):
        barrier(world)
        torch.cuda.synchronize()
        sampler.begin()
        t0 = time.perf_counter()
        losses_read = 0
        for j, (i, n) in enumerate(chunks):
            if n == spe:
                model.train_on_batches(xs[(rep + j) % 8], ys[(rep + j) % 8], steps=n, global_batch=gb)
            else:
                model.train_on_batches(xt[(rep + j) % 8], yt[(rep + j) % 8], steps=n, global_batch=gb)
            if j > 0:
                ls = model.last_losses(lag=1)      # the previous execution's losses, read while this one runs
                losses_read += len(ls)
        ls = model.last_losses()
        losses_read += len(ls)
        loss = ls[-1]
        assert losses_read == K, (losses_read, K)
        torch.cuda.synchronize()
        e2e_s.append(max_over_ranks(time.perf_counter() - t0, world))
        sampler.end()
    clocks = sampler.stop()
    barrier(world)
    e2e_med = statistics.median(e2e_s)

    if world > 1:
        strategy.comm().check_health()
    if fused:
        model.fused_plan().check()
        ar = ("in-kernel reduce-scatter + all-gather over peer memory inside the whole-step kernel "
              "(no NCCL, no separate launch)") if world > 1 else "n/a"
    else:
        ch = getattr(model.arena, "_channel", None)
        mode = {0: "p2p one-shot", 1: "p2p two-shot", 2: "NVLS multimem"}.get(getattr(ch, "mode", -1), "?")
        ar = f"fused {mode} all-reduce + optimizer kernel (no NCCL on step path)" if world > 1 else "n/a"
    standin = _standin_number(world, args)
    value = gb * K / (dev_med / 1e3)
    out = {
        "metric": METRIC, "value": value, "unit": "images/sec", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": dev_med / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "fp32" if args.dtype == "fp32" else "bf16 (mixed: fp32 master weights, tcgen05 dense GEMMs)",
        "data": "synthetic", "impl": "ours",
        "config": {"model": "keras_mnist_cnn_conv32_pool_dense64_dense10" if args.model == "a" else "model_b_bn_cnn",
                   "global_batch": gb, "per_gpu_batch": B, "seq_len": None, "input": list(in_shape),
                   "parallelism": f"dp{world}", "optimizer": args.optimizer,
                   "whole_step_kernel": fused, "steps_per_execution": spe,
                   "steps_per_execution_device_timed": spe_dev,
                   "l2": (f"inputs larger than L2: each timed step reads a distinct device-resident batch from a "
                          f"{(P * per) >> 20} MiB pool (L2 evicted by a 256 MiB write before each repeat); K steps "
                          f"back to back, one event pair; weights stay resident as in training"),
                   "repeats": R, "ms_per_repeat": dev_ms, "statistic": "median over repeats of max over ranks",
                   "allreduce": ar},
        "clocks": clocks,
        "e2e": {"value": gb * K / e2e_med, "unit": "images/sec", "ms_per_step": 1e3 * e2e_med / K,
                "h2d_bytes_per_step": per + (4 if (fused and _fs.streamed_inputs()) else 0), "d2h_bytes_per_step": 16, "last_loss": loss,
                "s_per_repeat": e2e_s,
                "note": ("model.train_on_batches(x, y, steps) == Keras steps_per_execution: per step one cudaMemcpyAsync "
                         "of that step's pinned uint8 images, one of its int64 labels and a 4-byte ready tag (copy "
                         "stream); one launch per execution, started without waiting for the copies -- the kernel takes "
                         "up step s when its tag has landed (TFDE_FUSED_STREAM=0: all copies first, then the launch); "
                         "every step's [loss, correct, rows, 0] written into mapped pinned memory and read on the host "
                         "(K reads for K steps)")},
        "gpu_launches": int(launches),
        "kernel_launches_per_step": launches / K,
        "correctness": check,
        "vs_standin": standin,
    }
    if standin and standin.get("value"):
        out["vs_standin"]["ratio"] = value / standin["value"]
        if standin.get("e2e"):
            out["vs_standin"]["e2e_ratio"] = out["e2e"]["value"] / standin["e2e"]
    if rank == 0:
        print(json.dumps(out), flush=True)
    strategy.shutdown()
    if world > 1 and torch.distributed.is_initialized():
        torch.distributed.destroy_process_group()


def _standin_number(world, args):
    """Recorded measurement of the CUDA-graphed torch/cuDNN/cuBLAS/NCCL stand-in on this pool (bench.py --impl torch,
    baseline/standin_results.json). NOT the reference (TensorFlow cannot be installed here): `vs_baseline` stays
    null; this ratio is reported under its own name."""
    if args.model != "a" or args.batch_per_gpu != 64 or args.dtype != "fp32" or args.optimizer != "sgd":
        return None
    try:
        with open(os.path.join(ROOT, "baseline", "standin_results.json")) as f:
            rec = json.load(f).get(str(world))
        return dict(rec) if rec else None
    except Exception:
        return None


def run_reference(args):
    """The reference is three scripts that ``import tensorflow``; there is no TensorFlow wheel in
    /opt/wheelhouse and no network, and the reference has no setup.py/pyproject to install (DESIGN.md
    records the pip attempt). Nothing of this framework may stand in for it on this arm."""
    rank, _, _ = dist_env()
    why = "tensorflow is not installed and cannot be installed offline (no wheel in /opt/wheelhouse; reference has no installer)"
    try:
        import tensorflow  # noqa: F401
        why = "tensorflow import unexpectedly succeeded but no runner is wired; see baseline/"
    except Exception:
        pass
    if rank == 0:
        print(json.dumps({"impl": "reference", "unavailable": why}), flush=True)


def run_torch(args):
    from baseline import torch_ddp
    fn = torch_ddp.run if args.no_graph else torch_ddp.run_graphed
    fn(args, METRIC, ClockSampler, max_over_ranks, barrier, dist_env, make_batches)


def main():
    args = parse_args()
    if args.impl == "reference":
        return run_reference(args)
    if args.impl == "torch":
        return run_torch(args)
    return run_ours(args)


if __name__ == "__main__":
    main()
