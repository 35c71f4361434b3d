"""ResNet-50 (v1, Keras ``applications`` topology) for BASELINE.json config 4: "ResNet-50 synthetic 224x224
under the same strategy API, to show scaling beyond a latency-bound toy model". Not part of the reference
(SURVEY.md section 8 row 18); it exists to exercise the strategy, the 102 MB two-shot fused all-reduce and
the conv / BN / pooling kernels at a scale that is not launch-bound.

The network is a ``Sequential`` whose residual stages are :class:`Bottleneck` composite layers (main path
1x1 -> 3x3 -> 1x1 with BN(+ReLU), identity or projection shortcut, add + ReLU), each with an explicit
backward that sends the gradient down both branches and sums it.
"""
from __future__ import annotations

from typing import List

import torch

from .. import ops
from ..keras import layers as L
from ..keras.models import Sequential


class Bottleneck(L.Layer):
    _type_name = "bottleneck"

    def __init__(self, filters: int, stride: int = 1, projection: bool = False, expansion: int = 4, **kw):
        super().__init__(**kw)
        self.filters, self.stride, self.projection, self.expansion = int(filters), int(stride), bool(projection), int(expansion)
        self.subs: List[L.Layer] = []

    def _mk(self, layer, suffix):
        layer.name = f"{self.name}_{suffix}"
        layer.model = self.model
        layer.need_dx = True
        self.subs.append(layer)
        return layer

    def build(self, in_shape):
        f, e = self.filters, self.expansion
        self.subs = []
        self.c1 = self._mk(L.Conv2D(f, 1, strides=self.stride), "1_conv")
        self.b1 = self._mk(L.BatchNormalization(epsilon=1.001e-5), "1_bn")
        self.c2 = self._mk(L.Conv2D(f, 3, padding="same"), "2_conv")
        self.b2 = self._mk(L.BatchNormalization(epsilon=1.001e-5), "2_bn")
        self.c3 = self._mk(L.Conv2D(f * e, 1), "3_conv")
        self.b3 = self._mk(L.BatchNormalization(epsilon=1.001e-5), "3_bn")
        self.b1.fused_relu = self.b2.fused_relu = True
        self.c1.bias_grad_zero = self.c2.bias_grad_zero = self.c3.bias_grad_zero = True   # each feeds a batch-norm
        for conv, bn in ((self.c1, self.b1), (self.c2, self.b2), (self.c3, self.b3)):
            conv.feeds_bn, bn.stats_source = True, conv       # BN statistics come out of the conv GEMM's epilogue
        shape = tuple(in_shape)
        for l in (self.c1, self.b1, self.c2, self.b2, self.c3, self.b3):
            shape = l.build(shape)
        if self.projection:
            self.c0 = self._mk(L.Conv2D(f * e, 1, strides=self.stride), "0_conv")
            self.b0 = self._mk(L.BatchNormalization(epsilon=1.001e-5), "0_bn")
            self.c0.bias_grad_zero = True
            self.c0.feeds_bn, self.b0.stats_source = True, self.c0
            s0 = self.b0.build(self.c0.build(tuple(in_shape)))
            if s0 != shape:
                raise ValueError(f"{self.name}: shortcut {s0} != main {shape}")
        elif tuple(in_shape) != shape:
            raise ValueError(f"{self.name}: identity shortcut needs matching shapes ({in_shape} vs {shape})")
        self.variables = [v for l in self.subs for v in l.variables]
        self.in_shape, self.out_shape, self.built = tuple(in_shape), shape, True
        return shape

    def takes_bf16(self, x) -> bool:
        return x.shape[-1] % 8 == 0

    def forward(self, x, training):
        for l in self.subs:
            l.model = self.model
        h = x
        for l in (self.c1, self.b1, self.c2, self.b2, self.c3, self.b3):
            h = l.run(h, training)
        sc = self.b0.run(self.c0.run(x, training), training) if self.projection else x
        self._b16 = h.dtype == torch.bfloat16 and sc.dtype == torch.bfloat16
        if self._b16:       # bf16 activations end to end (ops/act16.py)
            from ..ops import act16
            y = self.buf("y16", h.shape, torch.bfloat16)
            act16.add(h, sc, relu=True, out=y)
        else:
            if h.dtype != torch.float32 or sc.dtype != torch.float32:
                h, sc = h.float(), sc.float()
            y = self.buf("y", h.shape)
            ops.add(h, sc, relu=True, out=y)
        self._y = y
        return y

    def backward(self, dy):
        if self._b16:
            from ..ops import act16
            d = act16.relu_bwd(self._y, dy, out=self.buf("d16", dy.shape, torch.bfloat16))
        else:
            d = ops.relu_bwd(self._y, dy, out=self.buf("d", dy.shape))
        g = d
        for l in (self.b3, self.c3, self.b2, self.c2, self.b1, self.c1):
            g = l.run_back(g)
        gs = self.c0.run_back(self.b0.run_back(d)) if self.projection else d
        if g.dtype == torch.bfloat16 and gs.dtype == torch.bfloat16:
            from ..ops import act16
            return act16.add(g, gs, relu=False, out=self.buf("dx16", g.shape, torch.bfloat16))
        if g.dtype != torch.float32 or gs.dtype != torch.float32:
            g, gs = g.float(), gs.float()
        return ops.add(g, gs, relu=False, out=self.buf("dx", g.shape))

    def free_buffers(self):
        super().free_buffers()
        for l in self.subs:
            l.free_buffers()


def build_resnet(blocks=(3, 4, 6, 3), width=64, classes=1000, input_shape=(224, 224, 3), seed=None,
                 name="resnet50") -> Sequential:
    layers: List[L.Layer] = [
        L.Conv2D(width, 7, strides=2, padding="same", use_bias=True, input_shape=input_shape, name="conv1_conv"),
        L.BatchNormalization(epsilon=1.001e-5, name="conv1_bn"),
        L.Activation("relu", name="conv1_relu"),
        L.MaxPooling2D(3, strides=2, padding="same", name="pool1_pool"),
    ]
    f = width
    for stage, n in enumerate(blocks):
        for b in range(n):
            stride = 2 if (b == 0 and stage > 0) else 1
            layers.append(Bottleneck(f, stride=stride, projection=(b == 0), name=f"conv{stage + 2}_block{b + 1}"))
        f *= 2
    layers += [L.GlobalAveragePooling2D(name="avg_pool"), L.Dense(classes, name="predictions")]
    return Sequential(layers, name=name, seed=seed)


def build_resnet50(classes=1000, input_shape=(224, 224, 3), seed=None) -> Sequential:
    """25.6 M parameters (25,583,592 trainable + 53,120 BN moving statistics with 1000 classes)."""
    return build_resnet((3, 4, 6, 3), 64, classes, input_shape, seed, "resnet50")


def build_and_compile_resnet50(optimizer=None, classes=1000, input_shape=(224, 224, 3), seed=None) -> Sequential:
    from ..keras import losses, optimizers
    m = build_resnet50(classes, input_shape, seed)
    m.compile(optimizer=optimizer or optimizers.SGD(0.1, momentum=0.9),
              loss=losses.SparseCategoricalCrossentropy(from_logits=True), metrics=["accuracy"])
    return m
