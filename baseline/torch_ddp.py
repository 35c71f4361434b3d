"""Stand-in baselines: the SAME model/config as bench.py run through PyTorch
+ cuDNN/cuBLAS + NCCL. Two flavours:

* ``run_graphed`` (default of ``bench.py --impl torch``): the WHOLE step -- uint8->float scaling, forward, loss,
  backward, one NCCL all-reduce of a flat gradient bucket, SGD -- captured in ONE CUDA graph with static inputs
  (cudnn.benchmark on, channels-last). This is the fair vendor-library denominator: no Python and no launch
  overhead between the ~25 library kernels of a step.
* ``run`` (``--no-graph``): eager + DistributedDataParallel/NCCL -- "a path that only calls NCCL for the gradient
all-reduce" (BASELINE.json). It exists because the TensorFlow reference cannot be installed on this
image; it is NOT the reference arm (``--impl reference`` reports unavailable) and none of the product's
kernels are used here. Built fairly: channels-last, cudnn.benchmark on, no host syncs inside a step.

Timing protocol is identical to the product arm: per-step CUDA events with an L2 flush between steps,
max over ranks; e2e region with pinned H2D input copy and a D2H loss read every step.
"""
from __future__ import annotations

import json
import os
import time


def build_model_a(torch):
    nn = torch.nn
    return nn.Sequential(nn.Conv2d(1, 32, 3), nn.ReLU(), nn.MaxPool2d(2), nn.Flatten(),
                         nn.Linear(5408, 64), nn.ReLU(), nn.Linear(64, 10))


def build_model_b(torch):
    """Model B of mnist_keras_distributed.py:79-109 in torch (NCHW, cuDNN): 3 x (Conv-BN-ReLU), Dense200-BN-ReLU,
    Dropout, Dense10. BatchNorm without scale (affine weight frozen at 1), eps 1e-3, momentum 0.01 (= Keras 0.99)."""
    nn = torch.nn

    def bn(c, two_d=True):
        m = (nn.BatchNorm2d if two_d else nn.BatchNorm1d)(c, eps=1e-3, momentum=0.01)
        m.weight.requires_grad_(False)
        return m
    return nn.Sequential(nn.Unflatten(1, (1, 28, 28)),
                         nn.Conv2d(1, 6, 3, padding=1, bias=False), bn(6), nn.ReLU(),
                         nn.Conv2d(6, 12, 6, stride=2, padding=2, bias=False), bn(12), nn.ReLU(),
                         nn.Conv2d(12, 24, 6, stride=2, padding=2, bias=False), bn(24), nn.ReLU(),
                         nn.Flatten(), nn.Linear(1176, 200, bias=False), bn(200, False), nn.ReLU(), nn.Dropout(0.5),
                         nn.Linear(200, 10))


def run(args, METRIC, ClockSampler, max_over_ranks, barrier, dist_env, make_batches):
    import torch
    import torch.distributed as dist
    rank, world, local = dist_env()
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1 and not dist.is_initialized():
        dist.init_process_group("cpu:gloo,cuda:nccl", device_id=dev)
    torch.backends.cudnn.benchmark = True
    torch.manual_seed(1234)
    model = build_model_a(torch).to(dev).to(memory_format=torch.channels_last)
    net = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local]) if world > 1 else model
    opt = torch.optim.SGD(net.parameters(), lr=0.001)
    lossf = torch.nn.CrossEntropyLoss()
    B = args.batch_per_gpu
    gb = B * world
    xs, ys = make_batches(8, B, seed=100 + rank)
    xs = [x.permute(0, 3, 1, 2).contiguous().pin_memory() for x in xs]
    in_x = torch.empty((B, 1, 28, 28), dtype=torch.uint8, device=dev)
    in_y = torch.empty((B,), dtype=torch.int64, device=dev)

    def step(x_u8, y):
        x = x_u8.float().div_(255.0).contiguous(memory_format=torch.channels_last)
        opt.zero_grad(set_to_none=True)
        loss = lossf(net(x), y)
        loss.backward()
        opt.step()
        return loss

    W = max(args.warmup, 3)
    for i in range(W):
        in_x.copy_(xs[i % 8], non_blocking=True); in_y.copy_(ys[i % 8], non_blocking=True)
        step(in_x, in_y)
    torch.cuda.synchronize()
    K = args.steps
    sampler = ClockSampler(local)
    sampler.start()
    time.sleep(0.05)

    def timed_flush():
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
        starts = [torch.cuda.Event(enable_timing=True) for _ in range(K)]
        ends = [torch.cuda.Event(enable_timing=True) for _ in range(K)]
        barrier(world); torch.cuda.synchronize(); sampler.begin()
        for i in range(K):
            flush.fill_(i & 0xFF)
            starts[i].record()
            step(in_x, in_y)
            ends[i].record()
        torch.cuda.synchronize(); sampler.end(); barrier(world)
        return max_over_ranks(sum(s.elapsed_time(e) for s, e in zip(starts, ends)), world)

    def timed_inputs(evict=True):
        # same protocol as bench.py run_ours: distinct device-resident batches from a pool larger than L2
        per = in_x.numel() + in_y.numel() * 8
        P = max(K, -(-(136 << 20) // per)) if evict else 8
        gen = torch.Generator(device=dev).manual_seed(4321 + rank)
        pool_x = torch.randint(0, 256, (P,) + tuple(in_x.shape), generator=gen, dtype=torch.uint8, device=dev)
        pool_y = torch.randint(0, 10, (P,) + tuple(in_y.shape), generator=gen, dtype=torch.int64, device=dev)
        if evict:
            torch.empty(256 << 20, dtype=torch.uint8, device=dev).fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier(world); torch.cuda.synchronize(); sampler.begin()
        e0.record()
        for i in range(K):
            step(pool_x[i % P], pool_y[i % P])
        e1.record()
        torch.cuda.synchronize(); sampler.end(); barrier(world)
        return max_over_ranks(e0.elapsed_time(e1), world)

    cold_ms = None
    if args.no_flush:
        dev_ms, l2_note = timed_inputs(evict=False), "not flushed (--no-flush)"
    elif getattr(args, "l2", "inputs") == "flush":
        dev_ms, l2_note = timed_flush(), "flushed between timed steps"
    else:
        dev_ms, l2_note = timed_inputs(), "inputs larger than L2 (distinct device-resident batch per step, K steps back to back)"
        cold_ms = timed_flush()
    barrier(world); torch.cuda.synchronize(); sampler.begin()
    t0 = time.perf_counter()
    lv = 0.0
    for i in range(K):
        in_x.copy_(xs[i % 8], non_blocking=True); in_y.copy_(ys[i % 8], non_blocking=True)
        lv = step(in_x, in_y).item()
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0, world)
    sampler.end()
    clocks = sampler.stop()
    h2d = xs[0].numel() + ys[0].numel() * 8
    out = {"metric": METRIC, "value": gb * K / (dev_ms / 1e3), "unit": "images/sec", "n_gpus": world, "steps": K,
           "warmup": W, "ms_per_step": dev_ms / K, "higher_is_better": True, "scaling": "weak",
           "vs_baseline": None, "dtype": "fp32", "data": "synthetic", "impl": "torch_eager_ddp_nccl_standin",
           "config": {"model": "keras_mnist_cnn_conv32_pool_dense64_dense10", "global_batch": gb,
                      "per_gpu_batch": B, "parallelism": f"dp{world}", "optimizer": "sgd",
                      "l2": l2_note},
           "clocks": clocks,
           "cold_l2": ({"value": gb * K / (cold_ms / 1e3), "ms_per_step": cold_ms / K} if cold_ms is not None else None),
           "e2e": {"value": gb * K / e2e_s, "unit": "images/sec", "ms_per_step": 1e3 * e2e_s / K,
                   "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4, "last_loss": lv},
           "gpu_launches": 0}
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_graphed(args, METRIC, ClockSampler, max_over_ranks, barrier, dist_env, make_batches):
    """Whole-step CUDA graph of the torch/cuDNN/cuBLAS/NCCL stand-in (see the module docstring)."""
    import statistics
    import torch
    import torch.distributed as dist
    rank, world, local = dist_env()
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1 and not dist.is_initialized():
        dist.init_process_group("cpu:gloo,cuda:nccl", device_id=dev)
    torch.backends.cudnn.benchmark = True
    torch.manual_seed(1234)
    is_b = getattr(args, "model", "a") == "b"
    model = (build_model_b(torch) if is_b else build_model_a(torch)).to(dev).to(memory_format=torch.channels_last)
    model.train()
    params = [p for p in model.parameters() if p.requires_grad]
    if world > 1:
        for p in model.parameters():
            dist.broadcast(p.data, src=0)
    # one flat gradient bucket: a single NCCL all-reduce per step, gradients are views into it
    flat = torch.zeros(sum(p.numel() for p in params), device=dev)
    off = 0
    for p in params:
        p.grad = flat[off:off + p.numel()].view(p.shape)
        off += p.numel()
    opt = torch.optim.SGD(params, lr=0.001)
    lossf = torch.nn.CrossEntropyLoss(reduction="sum")
    B = args.batch_per_gpu
    gb = B * world
    K, W = args.steps, max(args.warmup, 3)
    xs, ys = make_batches(8, B, seed=100 + rank)
    if is_b:       # the reference feeds Model B float rows of 784 (mnist_keras_distributed.py:211)
        xs = [x.reshape(B, 784).float().div_(255.0).pin_memory() for x in xs]
        in_x = torch.empty((B, 784), dtype=torch.float32, device=dev)
    else:
        xs = [x.permute(0, 3, 1, 2).contiguous().pin_memory() for x in xs]
        in_x = torch.empty((B, 1, 28, 28), dtype=torch.uint8, device=dev)
    in_y = torch.empty((B,), dtype=torch.int64, device=dev)
    loss_dev = torch.zeros((), device=dev)
    loss_host = [torch.zeros(()).pin_memory() for _ in range(2)]

    def step():
        x = in_x if is_b else in_x.float().div_(255.0).contiguous(memory_format=torch.channels_last)
        flat.zero_()
        loss = lossf(model(x), in_y) / gb          # SUM over ranks of per-rank sums / global batch
        loss.backward()
        if world > 1:
            dist.all_reduce(flat)
        opt.step()
        loss_dev.copy_(loss.detach())

    side = torch.cuda.Stream()
    in_x.copy_(xs[0]); in_y.copy_(ys[0])
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(max(W, 11)):
            step()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    barrier(world)
    graphs = []
    for k in range(2):                             # two graphs: each ends with the D2H of the loss into its own pinned slot
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            step()
            loss_host[k].copy_(loss_dev, non_blocking=True)
        graphs.append(g)
    torch.cuda.synchronize()
    for i in range(W):
        graphs[i & 1].replay()
    torch.cuda.synchronize()
    sampler = ClockSampler(local)
    sampler.start()
    time.sleep(0.05)
    REPS = 5
    per = in_x.numel() * in_x.element_size() + in_y.numel() * 8
    P = max(K * REPS, -(-(136 << 20) // per))
    gen = torch.Generator(device=dev).manual_seed(4321 + rank)
    if is_b:
        pool_x = torch.rand((P,) + tuple(in_x.shape), generator=gen, dtype=torch.float32, device=dev)
    else:
        pool_x = torch.randint(0, 256, (P,) + tuple(in_x.shape), generator=gen, dtype=torch.uint8, device=dev)
    pool_y = torch.randint(0, 10, (P,) + tuple(in_y.shape), generator=gen, dtype=torch.int64, device=dev)
    torch.empty(256 << 20, dtype=torch.uint8, device=dev).fill_(1)
    dev_ms = []
    for rep in range(REPS):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier(world); torch.cuda.synchronize(); sampler.begin()
        e0.record()
        for i in range(K):
            j = (rep * K + i) % P
            in_x.copy_(pool_x[j]); in_y.copy_(pool_y[j])
            graphs[i & 1].replay()
        e1.record()
        torch.cuda.synchronize(); sampler.end(); barrier(world)
        dev_ms.append(max_over_ranks(e0.elapsed_time(e1), world))
    dev_med = statistics.median(dev_ms)
    # end to end: pinned H2D every step, loss D2H every step (read one step late, like the product arm)
    e2e = []
    lv = 0.0
    evs = [torch.cuda.Event(), torch.cuda.Event()]
    for rep in range(REPS):
        barrier(world); torch.cuda.synchronize(); sampler.begin()
        t0 = time.perf_counter()
        for i in range(K):
            in_x.copy_(xs[i % 8], non_blocking=True); in_y.copy_(ys[i % 8], non_blocking=True)
            graphs[i & 1].replay()
            evs[i & 1].record()
            if i > 0:
                evs[(i - 1) & 1].synchronize()
                lv = float(loss_host[(i - 1) & 1])
        evs[(K - 1) & 1].synchronize()
        lv = float(loss_host[(K - 1) & 1])
        torch.cuda.synchronize()
        e2e.append(max_over_ranks(time.perf_counter() - t0, world))
        sampler.end()
    e2e_s = statistics.median(e2e)
    clocks = sampler.stop()
    out = {"metric": METRIC, "value": gb * K / (dev_med / 1e3), "unit": "images/sec", "n_gpus": world, "steps": K,
           "warmup": W, "ms_per_step": dev_med / K, "higher_is_better": True, "scaling": "weak",
           "vs_baseline": None, "dtype": "fp32", "data": "synthetic", "impl": "torch_cudagraph_nccl_standin",
           "config": {"model": "model_b_bn_cnn" if is_b else "keras_mnist_cnn_conv32_pool_dense64_dense10",
                      "global_batch": gb,
                      "per_gpu_batch": B, "parallelism": f"dp{world}", "optimizer": "sgd", "cuda_graph": True,
                      "l2": "inputs larger than L2 (distinct device-resident batch per step)",
                      "allreduce": "one NCCL all-reduce of a flat gradient bucket inside the graph" if world > 1 else "n/a",
                      "repeats": REPS, "ms_per_rep": dev_ms},
           "clocks": clocks,
           "e2e": {"value": gb * K / e2e_s, "unit": "images/sec", "ms_per_step": 1e3 * e2e_s / K,
                   "h2d_bytes_per_step": per, "d2h_bytes_per_step": 4, "last_loss": lv},
           "gpu_launches": 0}
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()
