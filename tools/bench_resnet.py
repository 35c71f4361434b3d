"""ResNet-50 synthetic 224x224 training throughput (BASELINE.json config 4) through the strategy API.
    python tools/bench_resnet.py --batch 32 --steps 10 [--impl torch]
Launch with torchrun for N > 1. Prints one JSON line (images/sec, device-timed, max over ranks)."""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=4)
    ap.add_argument("--size", type=int, default=224)
    ap.add_argument("--impl", default="ours", choices=["ours", "torch"])
    ap.add_argument("--dtype", default="fp32", choices=["fp32", "bf16"])
    ap.add_argument("--acts", default="bfloat16", choices=["float32", "bfloat16"],
                    help="activation dtype under --dtype bf16 (ours): bf16 layer outputs (default) or fp32")
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    g = torch.Generator().manual_seed(rank)
    x = torch.rand(a.batch, a.size, a.size, 3, generator=g)
    y = torch.randint(0, 1000, (a.batch,), generator=g)
    if a.impl == "ours":
        import tfde_b200 as tf
        if a.dtype == "bf16":
            tf.keras.mixed_precision.set_global_policy("mixed_bfloat16", activations=a.acts)
        st = tf.distribute.MirroredStrategy() if world > 1 else tf.distribute.get_strategy()
        with st.scope():
            m = tf.models.build_and_compile_resnet50(seed=1, input_shape=(a.size, a.size, 3))
        xd, yd = x.pin_memory(), y.pin_memory()
        step = lambda: m.train_on_batch(xd, yd, global_batch=a.batch * world)
        sync = m.sync
        loss = m.last_step_loss
    else:
        # stand-in baseline: torchvision ResNet-50, channels_last, cuDNN autotune, (bf16 autocast), DDP/NCCL
        import torch.distributed as dist
        import torchvision
        local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(local)
        dev = torch.device("cuda", local)
        if world > 1 and not dist.is_initialized():
            dist.init_process_group("cpu:gloo,cuda:nccl", device_id=dev)
        torch.backends.cudnn.benchmark = True
        net = torchvision.models.resnet50(num_classes=1000).to(dev).to(memory_format=torch.channels_last)
        ddp = torch.nn.parallel.DistributedDataParallel(net, device_ids=[local]) if world > 1 else net
        opt = torch.optim.SGD(ddp.parameters(), lr=0.1, momentum=0.9)
        xd = x.permute(0, 3, 1, 2).contiguous().to(dev).contiguous(memory_format=torch.channels_last)
        yd = y.to(dev)
        last = [None]

        def step():
            opt.zero_grad(set_to_none=True)
            with torch.autocast("cuda", dtype=torch.bfloat16, enabled=a.dtype == "bf16"):
                l = torch.nn.functional.cross_entropy(ddp(xd), yd)
            l.backward()
            opt.step()
            last[0] = l

        class _M:
            _stream = torch.cuda.current_stream()
            graph_captured = False
        m = _M()
        sync = torch.cuda.synchronize
        loss = lambda: float(last[0].item())
    for _ in range(a.warmup):
        step()
    sync()
    t0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(m._stream):
        e0.record(m._stream)
    for _ in range(a.steps):
        step()
    with torch.cuda.stream(m._stream):
        e1.record(m._stream)
    sync()
    ms = e0.elapsed_time(e1) / a.steps
    if world > 1:
        import torch.distributed as dist
        t = torch.tensor([ms], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = t.item()
    if rank == 0:
        print(json.dumps({"metric": "resnet50_train_images_per_sec", "value": a.batch * world / (ms * 1e-3),
                          "ms_per_step": ms, "n_gpus": world, "per_gpu_batch": a.batch, "dtype": a.dtype, "acts": a.acts if a.dtype == "bf16" else "float32",
                          "impl": a.impl, "loss": loss(), "wall_s": time.perf_counter() - t0,
                          "graph": m.graph_captured}), flush=True)
    if a.impl == "ours":
        st.shutdown()


if __name__ == "__main__":
    main()
