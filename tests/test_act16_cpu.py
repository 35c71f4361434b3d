"""bf16-activation policy (mixed_bfloat16 + activations='bfloat16') on the CPU restatement: dtype plumbing
through the layers, fall-back casts at layers without a bf16 path, and closeness to the fp32 model."""
import torch

import tfde_b200 as tf
from tfde_b200.models import resnet
from tfde_b200.ops import act16
from tfde_b200.parallel.strategy import Strategy

BF = torch.bfloat16


def _train(policy, acts, steps=4):
    tf.keras.mixed_precision.set_global_policy(policy, activations=acts)
    try:
        with Strategy(device="cpu").scope():
            m = resnet.build_resnet((1, 1, 1, 1), 16, 10, (32, 32, 3), seed=3)
            m.compile(optimizer=tf.keras.optimizers.SGD(0.05),
                      loss=tf.keras.losses.SparseCategoricalCrossentropy(from_logits=True), metrics=["accuracy"])
        g = torch.Generator().manual_seed(0)
        x = torch.rand(16, 32, 32, 3, generator=g)
        y = torch.randint(0, 10, (16,), generator=g)
        losses = []
        for _ in range(steps):
            m.train_on_batch(x, y)
            losses.append(m.last_step_loss())
        return losses, m
    finally:
        tf.keras.mixed_precision.set_global_policy("float32")


def test_policy_api():
    mp = tf.keras.mixed_precision
    assert mp.activation_dtype() == "float32"
    mp.set_global_policy("mixed_bfloat16", activations="bfloat16")
    assert mp.global_policy() == "mixed_bfloat16" and mp.activation_dtype() == "bfloat16"
    mp.set_global_policy("mixed_bfloat16")
    assert mp.activation_dtype() == "float32"
    mp.set_global_policy("float32")
    try:
        mp.set_global_policy("float32", activations="bfloat16")
        raise AssertionError("bf16 activations without the mixed policy must be rejected")
    except ValueError:
        pass


def test_resnet_trunk_runs_in_bf16_and_tracks_fp32():
    l16, m = _train("mixed_bfloat16", "bfloat16")
    l32, _ = _train("float32", None)
    dts = [getattr(l, "_out_dtype", None) for l in m.exec_plan]
    assert dts[0] == BF and dts[3] == BF and dts[4] == BF        # stem conv, pool, first block
    assert dts[-1] == torch.float32 and dts[-2] == torch.float32  # GAP hands fp32 to the classifier
    assert abs(l16[0] - l32[0]) < 0.1, (l16, l32)
    assert l16[-1] < 0.5 * l16[0] and l32[-1] < 0.5 * l32[0]
    # layers that fell back to fp32 (too few rows for a tensor-core tile) cast at their boundary and still train
    assert any(getattr(l, "_out_dtype", None) == torch.float32 for l in m.exec_plan[5:8])


def test_conv_bias_grad_zero_flag_matches_true_gradient():
    """d(loss)/d(bias) of conv -> training-mode BN is zero up to rounding: the skipped reduction changes nothing."""
    g = torch.Generator().manual_seed(1)
    x = torch.randn(4, 8, 8, 16, generator=g, dtype=torch.float64)
    w = torch.randn(1, 1, 16, 32, generator=g, dtype=torch.float64, requires_grad=False)
    b = torch.zeros(32, dtype=torch.float64, requires_grad=True)
    y = torch.einsum("bhwc,cd->bhwd", x, w[0, 0]) + b
    mu, var = y.mean((0, 1, 2)), y.var((0, 1, 2), unbiased=False)
    z = torch.relu((y - mu) / torch.sqrt(var + 1e-3))
    (z * torch.randn(z.shape, generator=g, dtype=torch.float64)).sum().backward()
    assert b.grad.abs().max().item() < 1e-10


def test_ops_roundtrip_dtypes():
    g = torch.Generator().manual_seed(2)
    a = torch.randn(2, 6, 6, 16, generator=g).to(BF)
    assert act16.add(a, a, relu=True).dtype == BF
    assert act16.global_avgpool_fwd(a).dtype == torch.float32
    assert act16.global_avgpool_bwd(torch.randn(2, 16), tuple(a.shape)).dtype == BF
    y, idx = act16.maxpool_fwd(a, (2, 2), (2, 2), (0, 0, 0, 0))
    assert y.dtype == BF and idx.dtype == torch.uint8
    assert act16.channels_supported(64) and act16.channels_supported(2048) and not act16.channels_supported(24)
