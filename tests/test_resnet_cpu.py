"""ResNet building blocks on the CPU path: parameter count of ResNet-50 and autograd parity of a tiny
residual network (projection + identity bottlenecks, BN with gamma, 3x3/s2 same max-pool, GAP)."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

import tfde_b200 as tf
from tfde_b200.parallel.strategy import Strategy


def test_resnet50_parameter_count():
    with Strategy(device="cpu").scope():
        m = tf.models.build_resnet50(input_shape=(64, 64, 3), seed=0)
        m.build()
    train = sum(v.size for v in m.trainable_variables)
    frozen = sum(v.size for v in m.non_trainable_variables)
    assert train == 25_583_592 and frozen == 53_120
    assert m.output_shape == (1000,)


def test_tiny_resnet_gradients_match_autograd():
    with Strategy(device="cpu").scope():
        m = tf.models.build_resnet((1, 2), width=4, classes=5, input_shape=(16, 16, 3), seed=3)
        m.compile(optimizer=tf.keras.optimizers.SGD(0.0),
                  loss=tf.keras.losses.SparseCategoricalCrossentropy(from_logits=True), metrics=["accuracy"])
    g = torch.Generator().manual_seed(0)
    x = torch.rand(6, 16, 16, 3, generator=g)
    y = torch.randint(0, 5, (6,), generator=g)
    W = {k: torch.tensor(v, dtype=torch.float64, requires_grad=True) for k, v in m.named_weights().items()}
    m.train_on_batch(x, y)

    def conv(h, name, stride=1, same=False, bias=True):
        k = W[f"{name}/kernel"]
        hp = h.permute(0, 3, 1, 2)
        if same:
            kh = k.shape[0]
            _, lo, hi = tf.ops.same_padding(h.shape[1], kh, stride)
            hp = F.pad(hp, (lo, hi, lo, hi))
        out = F.conv2d(hp, k.permute(3, 2, 0, 1), W[f"{name}/bias"] if bias else None, stride=stride)
        return out.permute(0, 2, 3, 1)

    def bn(h, name, relu):
        mu = h.mean((0, 1, 2)); var = h.var((0, 1, 2), unbiased=False)
        out = (h - mu) / torch.sqrt(var + 1.001e-5) * W[f"{name}/gamma"] + W[f"{name}/beta"]
        return torch.relu(out) if relu else out

    def block(h, name, stride, proj):
        z = bn(conv(h, f"{name}_1_conv", stride), f"{name}_1_bn", True)
        z = bn(conv(z, f"{name}_2_conv", 1, same=True), f"{name}_2_bn", True)
        z = bn(conv(z, f"{name}_3_conv"), f"{name}_3_bn", False)
        sc = bn(conv(h, f"{name}_0_conv", stride), f"{name}_0_bn", False) if proj else h
        return torch.relu(z + sc)

    h = bn(conv(x.double(), "conv1_conv", 2, same=True, bias=True), "conv1_bn", True)
    hp = F.pad(h.permute(0, 3, 1, 2), (0, 1, 0, 1), value=float("-inf"))   # 8 -> 4 with k3 s2 'same': pad (0,1)
    h = F.max_pool2d(hp, 3, 2).permute(0, 2, 3, 1)
    h = block(h, "conv2_block1", 1, True)
    h = block(h, "conv3_block1", 2, True)
    h = block(h, "conv3_block2", 1, False)
    z = h.mean((1, 2)) @ W["predictions/kernel"] + W["predictions/bias"]
    loss = F.cross_entropy(z, y)
    loss.backward()
    assert m.last_step_loss() == pytest.approx(loss.item(), rel=1e-4)
    for v in m.trainable_variables:
        np.testing.assert_allclose(v.grad.numpy(), W[v.name].grad.numpy(), rtol=5e-3, atol=5e-6, err_msg=v.name)
