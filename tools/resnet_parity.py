"""ResNet-50 bf16 step-by-step losses with the implicit-GEMM convolutions on and off (same seed, same batch):
the first steps must agree to bf16 accuracy; later steps drift apart chaotically (batch-norm + SGD on one batch)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch


def run(igemm, steps, batch, size):
    import tfde_b200 as tf
    from tfde_b200.ops import conv_igemm
    conv_igemm.ENABLED = bool(igemm)
    tf.keras.mixed_precision.set_global_policy("mixed_bfloat16", activations="bfloat16")
    tf.set_seed(1)
    m = tf.models.build_and_compile_resnet50(seed=1, input_shape=(size, size, 3))
    g = torch.Generator().manual_seed(0)
    x = torch.rand(batch, size, size, 3, generator=g).pin_memory()
    y = torch.randint(0, 1000, (batch,), generator=g).pin_memory()
    out = []
    for _ in range(steps):
        m.train_on_batch(x, y)
        out.append(round(m.last_step_loss(), 4))
    return out


if __name__ == "__main__":
    steps, batch, size = 6, 32, 128
    a = run(1, steps, batch, size)
    b = run(0, steps, batch, size)
    print("igemm on :", a)
    print("igemm off:", b)
    print("first-step |diff|:", abs(a[0] - b[0]), " second-step |diff|:", abs(a[1] - b[1]))
