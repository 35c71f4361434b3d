#!/usr/bin/env python
"""Per-phase timing of the whole-step kernel (globaltimer stamps inside the kernel, last step of a K-step launch).
Usage (on a GPU box): python tools/fused_step_phases.py [--steps 8] > profiles/fused_step_phases.txt"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import tfde_b200 as tf  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--batch", type=int, default=64)
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    strategy = tf.distribute.MirroredStrategy() if world > 1 else tf.distribute.get_strategy()
    with strategy.scope():
        m = tf.models.build_and_compile_cnn_model(learning_rate=0.001, seed=1)
    B, K = a.batch, a.steps
    g = torch.Generator().manual_seed(0)
    x = torch.randint(0, 256, (K * B, 28, 28, 1), generator=g, dtype=torch.uint8).cuda()
    y = torch.randint(0, 10, (K * B,), generator=g).cuda()
    m.train_on_batches(x, y, steps=K)
    m.sync()
    plan = m.fused_plan()
    for trial in range(3):
        res = plan.profile(x, y, B, K, B * world)
    if strategy.rank == 0:
        print(f"whole-step kernel, {world} GPU(s), batch {B}/GPU, step {max(K - 2, 0)} of a {K}-step launch "
              f"(us: mean over CTAs / max over CTAs, rank 0)")
        for k, v in res.items():
            print(f"  {k:<44s} {v[0]:8.2f} {v[1]:8.2f}")
    if world > 1:
        import torch.distributed as dist
        strategy.shutdown()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
