"""Multi-GPU worker (launched by torchrun, one rank per GPU): validates the fused peer-to-peer
all-reduce + optimizer kernel against a gathered fp64 host reference, checks that replicas stay
bit-identical, that the CUDA-graph-captured distributed step equals the eager one, and prints an
all-reduce bandwidth sweep. Exit code 0 == all checks passed on every rank."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np
import torch
import torch.distributed as dist

import tfde_b200 as tf
from tfde_b200.keras import optimizers as O
from tfde_b200.ops import native, reference as ref
from tfde_b200.parallel.comm import Arena


def gather_cpu(t):
    parts = [None] * dist.get_world_size()
    dist.all_gather_object(parts, t.detach().cpu().numpy())
    return parts


def check_fused(strategy, n, opt_name, mode, steps=3, wire_bf16=False):
    os.environ["TFDE_AR_MODE"] = mode
    comm = strategy.comm()
    comm.wire_bf16 = wire_bf16
    rank, world = strategy.rank, strategy.world_size
    mk = {"sgd": lambda: O.SGD(0.1), "momentum": lambda: O.SGD(0.1, momentum=0.9),
          "adam": lambda: O.Adam(0.01)}[opt_name]
    opt = mk().bind(strategy.device, seed=1)
    arena = Arena(n, strategy.device, comm=comm)
    g = torch.Generator().manual_seed(1234)
    w0 = torch.randn(n, generator=g)
    arena.w[:n] = w0.to(strategy.device)
    comm.prepare(arena, opt)
    # fp64 reference on the host
    w_ref = w0.double().clone()
    s1 = torch.zeros(n, dtype=torch.float64)
    s2 = torch.zeros(n, dtype=torch.float64)
    ok = True
    for step in range(steps):
        gl = torch.Generator().manual_seed(100 * step + rank)
        grad = torch.randn(n, generator=gl)
        arena.g[:n] = grad.to(strategy.device)
        arena.g[n:n + 4] = torch.tensor([0.25 * (rank + 1), float(rank), 64.0, 0.0], device=strategy.device)
        comm.reduce_and_apply(arena, opt)
        opt.iterations += 1
        torch.cuda.synchronize()
        comm.check_health()
        grads = gather_cpu(grad)
        if wire_bf16:
            grads = [torch.from_numpy(x).bfloat16().float().numpy() for x in grads]
        gsum = torch.from_numpy(np.sum(np.stack(grads).astype(np.float64), axis=0))
        if opt_name == "sgd":
            w_ref = ref.sgd_update(w_ref, gsum, 0.1)
        elif opt_name == "momentum":
            w_ref, s1 = ref.momentum_update(w_ref, gsum, s1, 0.1, 0.9)
        else:
            w_ref, s1, s2 = ref.adam_update(w_ref, gsum, s1, s2, 0.01, 0.9, 0.999, 1e-7, step + 1)
        got = arena.w[:n].cpu().double()
        err = (got - w_ref).abs().max().item()
        tol = 2e-4 if opt_name != "adam" else 2e-3
        if wire_bf16:
            tol *= 4
        if not err <= tol * max(1.0, w_ref.abs().max().item()):
            print(f"[rank {rank}] FAIL n={n} opt={opt_name} mode={mode} step={step} err={err:.3e}", flush=True)
            ok = False
        tail = arena.tail.cpu().tolist()
        exp0 = sum(0.25 * (r + 1) for r in range(world))
        if abs(tail[0] - exp0) > 1e-5 or abs(tail[2] - 64.0 * world) > 1e-3:
            print(f"[rank {rank}] FAIL tail {tail} expected loss {exp0}", flush=True)
            ok = False
    # replicas bit-identical
    ws = gather_cpu(arena.w[:n])
    if not all(np.array_equal(ws[0], x) for x in ws[1:]):
        print(f"[rank {rank}] FAIL replicas differ n={n} opt={opt_name} mode={mode}", flush=True)
        ok = False
    m = arena.metrics.cpu().tolist()
    if abs(m[2] - 64.0 * world * steps) > 1e-2:
        print(f"[rank {rank}] FAIL metrics {m}", flush=True)
        ok = False
    if native.hparams_get(opt.hparams)["step"] != steps:
        print(f"[rank {rank}] FAIL device step {native.hparams_get(opt.hparams)['step']} != {steps}", flush=True)
        ok = False
    return ok


def check_model(strategy, graph):
    """2+ replicas, fixed data: weights identical across ranks and equal to a single-process CPU run on
    the concatenated (global) batch."""
    rank, world = strategy.rank, strategy.world_size
    os.environ.pop("TFDE_AR_MODE", None)
    with strategy.scope():
        m = tf.models.build_and_compile_cnn_model(learning_rate=0.05, seed=21)
    m.use_cuda_graph = graph
    g = torch.Generator().manual_seed(5)
    B = 32
    X = torch.rand(6, world * B, 28, 28, 1, generator=g)
    Y = torch.randint(0, 10, (6, world * B), generator=g)
    losses = []
    for s in range(6):
        m.train_on_batch(X[s, rank * B:(rank + 1) * B], Y[s, rank * B:(rank + 1) * B], global_batch=world * B)
        losses.append(m.last_step_loss())
    strategy.comm().check_health()
    from tfde_b200.parallel.strategy import Strategy
    with Strategy(device="cpu").scope():
        mc = tf.models.build_and_compile_cnn_model(learning_rate=0.05, seed=21)
    lc = []
    for s in range(6):
        mc.train_on_batch(X[s], Y[s])
        lc.append(mc.last_step_loss())
    ok = np.allclose(losses, lc, rtol=2e-3, atol=2e-4)
    if not ok:
        print(f"[rank {rank}] FAIL model losses {losses} vs cpu {lc}", flush=True)
    ws = gather_cpu(m.arena.w[:m.arena.n])
    same = all(np.array_equal(ws[0], x) for x in ws[1:])
    if not same:
        print(f"[rank {rank}] FAIL model replicas differ (graph={graph})", flush=True)
    close = all(np.allclose(a, b, rtol=5e-3, atol=2e-4) for a, b in zip(m.get_weights(), mc.get_weights()))
    if not close:
        print(f"[rank {rank}] FAIL model weights differ from cpu reference", flush=True)
    from tfde_b200.ops import push as _push
    nvls_mode = getattr(m.arena._channel, "mode", -1) == 2      # no receive slots to push into under NVLS
    if os.environ.get("TFDE_FUSED_PUSH", "1") == "1" and not nvls_mode and _push.stats["pushed_launches"] == 0:
        print(f"[rank {rank}] FAIL fused wgrad->push was never used", flush=True)
        ok = False
    if graph and not m.graph_captured:
        print(f"[rank {rank}] FAIL graph not captured", flush=True)
        ok = False
    return ok and same and close


def check_ps_push_pull(strategy):
    """G16: every rank pushes -lr*g into rank 0's shard with NVLink reduction atomics, then pulls it."""
    import ctypes
    L = native.load()
    L.tfde_ps_push.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_long, ctypes.c_float, ctypes.c_void_p]
    L.tfde_ps_pull.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_long, ctypes.c_void_p]
    comm = strategy.comm()
    rank, world = strategy.rank, strategy.world_size
    n = 100003
    shard = comm.symm_empty(n + 1, "ps_shard")        # owner copy lives on rank 0
    buf = comm.buffer_of(shard)
    w0 = torch.randn(n, generator=torch.Generator().manual_seed(9))
    if rank == 0:
        shard[:n] = w0.to(strategy.device)
    torch.cuda.synchronize(); dist.barrier()
    g = torch.randn(n, generator=torch.Generator().manual_seed(50 + rank)).to(strategy.device)
    native.check(L.tfde_ps_push(buf.peer_ptrs[0], g.data_ptr(), n, 0.1, native.stream_ptr()), "ps_push")
    torch.cuda.synchronize(); dist.barrier()
    local = torch.empty(n, device=strategy.device)
    native.check(L.tfde_ps_pull(local.data_ptr(), buf.peer_ptrs[0], n, native.stream_ptr()), "ps_pull")
    torch.cuda.synchronize()
    exp = w0.double()
    for r in range(world):
        exp = exp - 0.1 * torch.randn(n, generator=torch.Generator().manual_seed(50 + r)).double()
    err = (local.cpu().double() - exp).abs().max().item()
    ok = err < 1e-5
    if not ok:
        print(f"[rank {rank}] FAIL ps push/pull err {err:.3e}", flush=True)
    dist.barrier()
    return ok


def soak(strategy, iters):
    """Stress / soak of the cross-GPU protocol (what a single-device race checker cannot see): `iters`
    back-to-back fused all-reduce+SGD launches per variant with random per-rank host delays (ranks arrive at the
    flag barriers in every order), several arena sizes, integer-valued gradients so that every partial sum is
    exact in fp32 and the expected weights are known in closed form. Checks: exact result, bit-identical
    replicas, no timeout status, epochs in lock-step."""
    import random
    comm = strategy.comm()
    rank, world = strategy.rank, strategy.world_size
    rnd = random.Random(1000 + rank)
    ok = True
    modes = ("one_shot", "two_shot") + (("nvls",) if comm.nvls else ())
    for mode in modes:
        for n in (257, 40_003, 347_146):
            os.environ["TFDE_AR_MODE"] = mode
            opt = O.SGD(1.0).bind(strategy.device, seed=1)
            arena = Arena(n, strategy.device, comm=comm)
            comm.prepare(arena, opt)
            arena.w.zero_()
            base = (torch.arange(arena.n_total, device=strategy.device) % 7).float()
            for it in range(iters):
                arena.g.copy_(base * float(rank + 1))          # sum over ranks = base * W(W+1)/2
                if rnd.random() < 0.3:
                    time.sleep(rnd.random() * 2e-3)            # skew the ranks against each other
                comm.reduce_and_apply(arena, opt)
                opt.iterations += 1
                if it % 64 == 63:
                    torch.cuda.synchronize()
            torch.cuda.synchronize()
            comm.check_health()
            expect = -(base[:n].double() * (world * (world + 1) / 2) * iters)
            err = (arena.w[:n].double() - expect).abs().max().item()
            if err != 0.0:
                print(f"[rank {rank}] FAIL soak mode={mode} n={n}: max err {err}", flush=True)
                ok = False
            ws = gather_cpu(arena.w[:n])
            if not all(np.array_equal(ws[0], x) for x in ws[1:]):
                print(f"[rank {rank}] FAIL soak replicas differ mode={mode} n={n}", flush=True)
                ok = False
            if native.hparams_get(opt.hparams)["step"] != iters:
                print(f"[rank {rank}] FAIL soak device step counter mode={mode} n={n}", flush=True)
                ok = False
            del arena
    return ok


def sweep(strategy, out_path):
    """Fused all-reduce(+SGD) bus bandwidth: S/t * 2(N-1)/N, device-timed, max over ranks."""
    comm = strategy.comm()
    rank, world = strategy.rank, strategy.world_size
    rows = []
    modes = ("one_shot", "two_shot") + (("nvls",) if comm.nvls else ())
    if os.environ.get("TFDE_SWEEP_MODES"):
        modes = tuple(os.environ["TFDE_SWEEP_MODES"].split(","))
    for mode in modes:
        top = int(os.environ.get("TFDE_SWEEP_MAX", str(1 << 30)))
        for nbytes in (1 << 10, 1 << 14, 1 << 18, 1388584, 1 << 22, 1 << 24, 1 << 26, 1 << 28, 1 << 30):
            if nbytes > top or (mode == "one_shot" and nbytes * world > (1 << 28)):
                continue
            n = nbytes // 4
            os.environ["TFDE_AR_MODE"] = mode
            opt = O.SGD(0.0).bind(strategy.device)
            arena = Arena(n, strategy.device, comm=comm)
            comm.prepare(arena, opt)
            arena.g.normal_()
            iters = 20 if nbytes <= (1 << 26) else 5

            def timed(fn):
                """CUDA-graph replay of `iters` back-to-back launches (removes host launch overhead);
                falls back to eager launches if capture is not possible. Device time, max over ranks."""
                stream = torch.cuda.Stream()
                with torch.cuda.stream(stream):
                    for _ in range(3):
                        fn()
                stream.synchronize(); dist.barrier()
                graph = None
                try:
                    graph = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(graph, stream=stream):
                        for _ in range(iters):
                            fn()
                except Exception as ex:  # noqa: BLE001
                    if rank == 0:
                        print("graph capture failed, timing eager launches:", ex, flush=True)
                    graph = None
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                with torch.cuda.stream(stream):
                    if graph is not None:
                        graph.replay()
                    stream.synchronize(); dist.barrier()
                    e0.record(stream)
                    if graph is not None:
                        graph.replay()
                    else:
                        for _ in range(iters):
                            fn()
                    e1.record(stream)
                stream.synchronize()
                tt = torch.tensor([e0.elapsed_time(e1) / iters], dtype=torch.float64)
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
                return tt.item()

            ms = timed(lambda: comm.reduce_and_apply(arena, opt))
            buf = torch.randn(n, device=strategy.device)
            nccl_ms = timed(lambda: dist.all_reduce(buf))
            bus = lambda ms_: nbytes / (ms_ * 1e-3) * 2 * (world - 1) / world / 1e9
            rows.append({"mode": mode, "bytes": nbytes, "world": world, "fused_us": ms * 1e3,
                         "fused_bus_GBps": bus(ms), "nccl_allreduce_only_us": nccl_ms * 1e3,
                         "nccl_bus_GBps": bus(nccl_ms)})
            comm.check_health()
            del arena, buf
    if rank == 0:
        with open(out_path, "w") as f:
            json.dump(rows, f, indent=1)
        for r in rows:
            print(f"{r['mode']:9s} {r['bytes']:>10d} B  fused(all-reduce+SGD) {r['fused_us']:9.1f} us "
                  f"{r['fused_bus_GBps']:7.1f} GB/s | nccl all-reduce only {r['nccl_allreduce_only_us']:9.1f} us "
                  f"{r['nccl_bus_GBps']:7.1f} GB/s", flush=True)


def main():
    strategy = tf.distribute.MirroredStrategy()
    rank, world = strategy.rank, strategy.world_size
    assert world >= 2, "launch with torchrun --nproc-per-node >= 2"
    ok = True
    if "--sweep-only" in sys.argv:
        sweep(strategy, os.path.join(ROOT, "gpurun_out", os.environ.get("TFDE_SWEEP_OUT", f"allreduce_sweep_n{world}.json")))
        strategy.shutdown()
        dist.destroy_process_group()
        return
    if "--soak" in sys.argv:
        iters = int(sys.argv[sys.argv.index("--soak") + 1])
        sok = soak(strategy, iters)
        flag = torch.tensor([1.0 if sok else 0.0], dtype=torch.float64)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if rank == 0:
            print(f"P2P SOAK ({iters} iterations x 3 sizes x variants)", "PASSED" if flag.item() == 1.0 else "FAILED",
                  flush=True)
        strategy.shutdown()
        dist.destroy_process_group()
        sys.exit(0 if flag.item() == 1.0 else 1)
    if rank == 0:
        print("NVLS multicast:", "available" if strategy.comm().nvls else "unavailable", flush=True)
    # small scalar all-reduce (G14)
    t = torch.arange(7, dtype=torch.float32, device=strategy.device) + rank
    strategy.all_reduce_sum_(t)
    exp = torch.arange(7, dtype=torch.float32) * world + sum(range(world))
    if not torch.equal(t.cpu(), exp):
        print(f"[rank {rank}] FAIL scalar allreduce {t.cpu()} vs {exp}", flush=True)
        ok = False
    for n in (1000, 347146, 3_000_001):
        for opt_name in ("sgd", "momentum", "adam"):
            for mode in ("one_shot", "two_shot") + (("nvls",) if strategy.comm().nvls else ()):
                ok &= check_fused(strategy, n, opt_name, mode)
    ok &= check_fused(strategy, 347146, "sgd", "one_shot", wire_bf16=True)
    ok &= check_fused(strategy, 347146, "adam", "two_shot", wire_bf16=True)
    strategy.comm().wire_bf16 = False
    ok &= check_ps_push_pull(strategy)
    ok &= check_model(strategy, graph=False)
    ok &= check_model(strategy, graph=True)
    flag = torch.tensor([1.0 if ok else 0.0], dtype=torch.float64)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("P2P CHECKS", "PASSED" if flag.item() == 1.0 else "FAILED", flush=True)
    if "--sweep" in sys.argv and flag.item() == 1.0:
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        sweep(strategy, os.path.join(ROOT, "gpurun_out", f"allreduce_sweep_n{world}.json"))
    strategy.shutdown()
    dist.destroy_process_group()
    sys.exit(0 if flag.item() == 1.0 else 1)


if __name__ == "__main__":
    main()
