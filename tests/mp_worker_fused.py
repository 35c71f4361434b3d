"""Multi-GPU worker (torchrun, one rank per GPU) for the whole-step kernel's in-kernel all-reduce
(csrc/kernels/fused_step.cu): after K steps on different per-rank data every replica must hold bit-identical
weights, and they must match ONE process training on the concatenated global batch through the PyTorch CPU
restatement. `--soak N` adds N more steps in multi-step launches with randomly delayed ranks (flag protocol under
skew). Exit code 0 and "FUSED STEP CHECKS PASSED" == every check passed on every rank."""
import argparse
import hashlib
import os
import random
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np
import torch
import torch.distributed as dist

import tfde_b200 as tf
from tfde_b200.parallel.strategy import Strategy


def digest(model):
    h = hashlib.sha256()
    for w in model.get_weights():
        h.update(w.tobytes())
    return h.hexdigest()


def all_same(strategy, value):
    vals = [None] * strategy.world_size
    dist.all_gather_object(vals, value)
    return all(v == vals[0] for v in vals)


def batches(rank, steps, B, seed):
    g = torch.Generator().manual_seed(seed + rank)
    x = torch.randint(0, 256, (steps, B, 28, 28, 1), generator=g, dtype=torch.uint8)
    y = torch.randint(0, 10, (steps, B), generator=g, dtype=torch.int64)
    return x, y


def mk_opt(kind):
    return {"sgd": lambda: tf.keras.optimizers.SGD(0.05),
            "momentum": lambda: tf.keras.optimizers.SGD(0.05, momentum=0.9),
            "adam": lambda: tf.keras.optimizers.Adam(0.002)}[kind]()


def check(strategy, kind, B, steps=6):
    rank, world = strategy.rank, strategy.world_size
    gb = B * world
    with strategy.scope():
        m = tf.models.build_and_compile_cnn_model(optimizer=mk_opt(kind), seed=21)
    x, y = batches(rank, steps, B, 500)
    # first half step by step, second half as one multi-step launch
    half = steps // 2
    losses = []
    for k in range(half):
        m.train_on_batch(x[k], y[k], global_batch=gb)
        losses.append(m.last_step_loss())
    m.train_on_batches(x[half:], y[half:], global_batch=gb)
    losses += m.last_losses()
    m.sync()
    assert m.fused_plan() is not None, "whole-step kernel was not used"
    m.fused_plan().check()
    ok = all_same(strategy, digest(m)) and all_same(strategy, losses)
    err = 0.0
    if rank == 0:
        with Strategy(device="cpu").scope():
            mc = tf.models.build_and_compile_cnn_model(optimizer=mk_opt(kind), seed=21)
        data = [batches(r, steps, B, 500) for r in range(world)]
        lc = []
        for k in range(steps):
            mc.train_on_batch(torch.cat([d[0][k] for d in data], 0), torch.cat([d[1][k] for d in data], 0), global_batch=gb)
            lc.append(mc.last_step_loss())
        for a, b in zip(m.get_weights(), mc.get_weights()):
            err = max(err, float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-12)))
        lerr = float(np.abs(np.array(losses) - np.array(lc)).max())
        ok = ok and err < 2e-3 and lerr < 2e-4
        print(f"[fused {kind} B={B} W={world}] replicas identical={ok} max_rel_err_vs_cpu={err:.3e} loss_err={lerr:.3e}",
              flush=True)
    return ok


def soak(strategy, n_steps, B=64):
    rank, world = strategy.rank, strategy.world_size
    gb = B * world
    with strategy.scope():
        m = tf.models.build_and_compile_cnn_model(optimizer=mk_opt("sgd"), seed=5)
    rng = random.Random(1000 + rank)      # per-rank: delays and host/device input choice
    shared = random.Random(77)            # same on every rank: steps per launch
    done = 0
    x, y = batches(rank, 16, B, 900)
    xd, yd = x.cuda(), y.cuda()
    while done < n_steps:
        k = shared.choice([1, 2, 5, 16])
        if rng.random() < 0.5:
            torch.cuda._sleep(int(rng.random() * 2e6))       # skew this rank by up to ~1 ms
        if rng.random() < 0.5:
            m.train_on_batches(xd[:k], yd[:k], global_batch=gb)
        else:
            m.train_on_batches(x[:k], y[:k], global_batch=gb)
        done += k
    m.sync()
    m.fused_plan().check()
    ok = all_same(strategy, digest(m))
    if rank == 0:
        print(f"[fused soak W={world}] {done} steps, replicas identical={ok}", flush=True)
    return ok


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--soak", type=int, default=0)
    a = ap.parse_args()
    strategy = tf.distribute.MirroredStrategy()
    ok = True
    for kind, B in [("sgd", 64), ("adam", 64), ("momentum", 37), ("sgd", 1)]:
        ok = check(strategy, kind, B) and ok
    if a.soak:
        ok = soak(strategy, a.soak) and ok
    oks = [None] * strategy.world_size
    dist.all_gather_object(oks, bool(ok))
    if strategy.rank == 0:
        print("FUSED STEP CHECKS PASSED" if all(oks) else f"FUSED STEP CHECKS FAILED {oks}", flush=True)
    strategy.shutdown()
    dist.destroy_process_group()
    sys.exit(0 if all(oks) else 1)


if __name__ == "__main__":
    main()
