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

* ``value`` -- device-timed: every step (forward, loss, backward, fused cross-replica all-reduce +
  SGD update, metric accumulation -- one CUDA-graph launch) is bracketed by its own pair of CUDA events
  with an L2 flush (256 MiB write) *between* steps, outside the event pairs; per-rank total = sum of the
  K step times, job time = max over ranks.
* ``e2e`` -- wall clock through the public API (``model.train_on_batch``) including, every step, the
  host->device copy of that step's uint8 images + int64 labels from pinned memory and a device->host
  read of the step loss.

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

    def __init__(self, gpu_index=0, period_s=0.004):
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
    if args.dtype == "bf16":
        tf.keras.mixed_precision.set_global_policy("mixed_bfloat16")
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
    xs, ys = make_batches(8, B, seed=100 + rank)
    if args.model == "b":
        xs = [x.reshape(B, 784).float().div_(255.0).pin_memory() for x in xs]

    sampler = ClockSampler(local)
    sampler.start()        # started before the warm-up so that its own start-up is outside the timed regions
    # ---- warm-up (>= 3): eager steps, graph capture, first replays ---------------------------------
    W = max(args.warmup, 3)
    c0 = native.launch_count()
    for i in range(W):
        model.train_on_batch(xs[i % 8], ys[i % 8], global_batch=gb)
        if i == 0:
            model.sync()
            launches_per_step = native.launch_count() - c0
    model.sync()

    # ---- device-timed region -----------------------------------------------------------------------
    # Two protocols (both allowed by the timing rules; `config.l2` says which one `value` comes from):
    #  inputs: K back-to-back steps, one event pair around them; every step consumes a DISTINCT device-side
    #          input batch out of a pool larger than L2 (>= 136 MiB, evicted once by a 256 MiB write before
    #          timing), so inputs always come from HBM while weights/activations are as resident as they are
    #          in real training. This is the default: it is what a training loop sees.
    #  flush:  a 256 MiB write between steps (outside per-step event pairs): everything cold every step.
    K = args.steps
    key = [k for k, v in model._graphs.items() if hasattr(v, "in_x")][0]
    st = model._graphs[key]
    st.in_x.copy_(xs[0]); st.in_y.copy_(ys[0])
    stream = model._stream
    barrier(world)
    torch.cuda.synchronize()
    with torch.cuda.stream(stream):
        step_graph = model.device_step_graph(st, gb) if model.use_cuda_graph else None

    def run_step():
        if step_graph is not None:
            step_graph.replay()
        else:
            model._step_body(st.in_x, st.in_y, gb)

    def timed_flush():
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
        starts = [torch.cuda.Event(enable_timing=True) for _ in range(K)]
        ends = [torch.cuda.Event(enable_timing=True) for _ in range(K)]
        barrier(world)
        torch.cuda.synchronize()
        sampler.begin()
        with torch.cuda.stream(stream):
            for i in range(K):
                flush.fill_(i & 0xFF)
                starts[i].record(stream)
                run_step()
                ends[i].record(stream)
        torch.cuda.synchronize()
        sampler.end()
        barrier(world)
        return max_over_ranks(sum(s.elapsed_time(e) for s, e in zip(starts, ends)), world)

    def timed_inputs(evict=True):
        per = st.in_x.numel() * st.in_x.element_size() + st.in_y.numel() * 8
        P = max(K, -(-(136 << 20) // per)) if evict else 8
        gen = torch.Generator(device=dev).manual_seed(4321 + rank)
        if st.in_x.dtype == torch.uint8:
            pool_x = torch.randint(0, 256, (P,) + tuple(st.in_x.shape), generator=gen, dtype=torch.uint8, device=dev)
        else:
            pool_x = torch.rand((P,) + tuple(st.in_x.shape), generator=gen, dtype=st.in_x.dtype, device=dev)
        pool_y = torch.randint(0, 10, (P,) + tuple(st.in_y.shape), generator=gen, dtype=torch.int64, device=dev)
        if evict:
            torch.empty(256 << 20, dtype=torch.uint8, device=dev).fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier(world)
        torch.cuda.synchronize()
        sampler.begin()
        with torch.cuda.stream(stream):
            e0.record(stream)
            for i in range(K):
                st.in_x.copy_(pool_x[i % P]); st.in_y.copy_(pool_y[i % P])
                run_step()
            e1.record(stream)
        torch.cuda.synchronize()
        sampler.end()
        barrier(world)
        return max_over_ranks(e0.elapsed_time(e1), world), P * per

    cold_ms = None
    if args.no_flush:
        dev_ms, pool_bytes = timed_inputs(evict=False)
        l2_note = "not flushed, 8 input batches reused (--no-flush; not a valid headline number)"
    elif args.l2 == "flush":
        dev_ms = timed_flush()
        l2_note = "flushed between timed steps (256 MiB write outside the per-step event pairs)"
    else:
        dev_ms, pool_bytes = timed_inputs()
        l2_note = (f"inputs larger than L2: each timed step reads a distinct device-resident batch from a "
                   f"{pool_bytes >> 20} MiB pool (evicted by a 256 MiB write before timing); K steps back to back, "
                   f"one event pair; weights/activations stay resident as in training")
        cold_ms = timed_flush()
    model.optimizer.iterations += (K if step_graph is not None else 0) * (2 if cold_ms is not None else 1)

    # ---- end-to-end region (public API, H2D every step, D2H loss every step) -------------------------
    barrier(world)
    torch.cuda.synchronize()
    sampler.begin()
    t0 = time.perf_counter()
    loss = 0.0
    losses_read = 0
    for i in range(K):
        model.train_on_batch(xs[i % 8], ys[i % 8], global_batch=gb)
        if i > 0:
            loss = model.last_step_loss(lag=1)   # D2H result of step i-1, read while step i runs
            losses_read += 1
    loss = model.last_step_loss()                # ... and of the last step: K reads for K steps
    losses_read += 1
    assert losses_read == K
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0, world)
    sampler.end()
    clocks = sampler.stop()
    barrier(world)
    h2d = xs[0].numel() * xs[0].element_size() + ys[0].numel() * ys[0].element_size()

    ar_mode = "n/a"
    if world > 1:
        strategy.comm().check_health()
        ch = getattr(model.arena, "_channel", None)
        ar_mode = {0: "p2p one-shot", 1: "p2p two-shot", 2: "NVLS multimem"}.get(getattr(ch, "mode", -1), "?")
    ms_per_step = dev_ms / K
    value = gb * K / (dev_ms / 1e3)
    out = {
        "metric": METRIC, "value": value, "unit": "images/sec", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "fp32" if args.dtype == "fp32" else "bf16 (mixed: fp32 master weights, tcgen05 dense GEMMs)",
        "data": "synthetic", "impl": "ours",
        "config": {"model": "keras_mnist_cnn_conv32_pool_dense64_dense10" if args.model == "a" else "model_b_bn_cnn",
                   "global_batch": gb, "per_gpu_batch": B, "seq_len": None, "input": list(in_shape),
                   "parallelism": f"dp{world}", "optimizer": args.optimizer, "cuda_graph": model.graph_captured,
                   "l2": l2_note,
                   "allreduce": (f"fused {ar_mode} all-reduce + optimizer kernel (no NCCL on step path)"
                                 if world > 1 else "n/a")},
        "clocks": clocks,
        "cold_l2": ({"value": gb * K / (cold_ms / 1e3), "ms_per_step": cold_ms / K,
                     "note": "same K steps with a 256 MiB L2 flush before every step (per-step event pairs)"}
                    if cold_ms is not None else None),
        "e2e": {"value": gb * K / e2e_s, "unit": "images/sec", "ms_per_step": 1e3 * e2e_s / K,
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4, "last_loss": loss,
                "note": "model.train_on_batch + last_step_loss(lag=1): pinned uint8 images + int64 labels H2D and loss D2H every step; each step's loss is read on the host one step late (K reads for K steps) so the read overlaps the next step"},
        "gpu_launches": int(launches_per_step * K),
        "kernel_launches_per_step": int(launches_per_step),
    }
    if rank == 0:
        print(json.dumps(out), flush=True)
    strategy.shutdown()
    if world > 1 and torch.distributed.is_initialized():
        torch.distributed.destroy_process_group()


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
    from baseline.torch_ddp import run as run_baseline
    run_baseline(args, METRIC, ClockSampler, max_over_ranks, barrier, dist_env, make_batches)


def main():
    args = parse_args()
    if args.impl == "reference":
        return run_reference(args)
    if args.impl == "torch":
        return run_torch(args)
    return run_ours(args)


if __name__ == "__main__":
    main()
