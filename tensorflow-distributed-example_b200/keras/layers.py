"""Keras-like layers with explicit forward/backward (no autograd, no tracing).

Covers every layer the reference builds:
Conv2D / MaxPooling2D / Flatten / Dense (distributed_with_keras.py:33-39), Reshape /
BatchNormalization / Activation / Dropout (mnist_keras_distributed.py:79-109), plus Add /
GlobalAveragePooling2D / ZeroPadding-free padded pooling for the ResNet-50 config of BASELINE.json.

Layout is channels-last; variables use the Keras names, shapes and layouts (HWIO kernels, (in, out)
dense kernels, ``beta/gamma/moving_mean/moving_variance``) so checkpoints round-trip Keras weights
without transposes. Each layer saves what its backward needs in preallocated per-batch-size buffers
(static addresses -> the whole step is CUDA-graph capturable). Gradients are written straight into
views of the model's flat gradient arena, which is what the fused all-reduce consumes.
"""
from __future__ import annotations

import functools
import inspect
import math
from typing import List, Optional, Sequence, Tuple

import torch

from .. import ops

_ACTS = (None, "linear", "relu", "softmax")
_CACHE_IM2COL = __import__("os").environ.get("TFDE_CACHE_IM2COL", "1") == "1"
_FORK_CONV_WGRAD = __import__("os").environ.get("TFDE_FORK_CONV_WGRAD", "1") == "1"


def _pair(v) -> Tuple[int, int]:
    if isinstance(v, (tuple, list)):
        return int(v[0]), int(v[1])
    return int(v), int(v)


class Variable:
    """A named slice of one of the model's flat arenas."""

    def __init__(self, name, shape, initializer, trainable=True):
        self.name = name
        self.shape = tuple(int(s) for s in shape)
        self.initializer = initializer
        self.trainable = trainable
        self.value: Optional[torch.Tensor] = None  # view into params / state arena
        self.grad: Optional[torch.Tensor] = None   # view into grads arena (trainable only)
        self.offset = -1

    @property
    def size(self) -> int:
        n = 1
        for s in self.shape:
            n *= s
        return n

    def numpy(self):
        return self.value.detach().cpu().numpy().copy()


def glorot_uniform(shape, generator):
    """Keras default kernel initializer: U(-l, l), l = sqrt(6 / (fan_in + fan_out)); conv fans include
    the receptive field."""
    if len(shape) == 2:
        fan_in, fan_out = shape
    else:
        rf = 1
        for s in shape[:-2]:
            rf *= s
        fan_in, fan_out = shape[-2] * rf, shape[-1] * rf
    limit = math.sqrt(6.0 / (fan_in + fan_out))
    return (torch.rand(shape, generator=generator, dtype=torch.float32) * 2.0 - 1.0) * limit


def zeros_init(shape, generator):
    return torch.zeros(shape, dtype=torch.float32)


def ones_init(shape, generator):
    return torch.ones(shape, dtype=torch.float32)


class Layer:
    _type_name = "layer"
    has_relu_output = False  # set by fusion pass consumers

    def __init_subclass__(cls, **kw):
        """Records the constructor arguments of every layer so get_config()/layer_from_config() can
        rebuild it (used by the serving export)."""
        super().__init_subclass__(**kw)
        orig = cls.__init__
        if getattr(orig, "_records_config", False):
            return
        sig = inspect.signature(orig)

        @functools.wraps(orig)
        def init(self, *a, **k):
            if not hasattr(self, "_ctor_config"):
                try:
                    ba = sig.bind(self, *a, **k)
                    cfg = {}
                    for n, v in list(ba.arguments.items())[1:]:
                        if sig.parameters[n].kind is inspect.Parameter.VAR_KEYWORD:
                            cfg.update(v)
                        else:
                            cfg[n] = v
                    self._ctor_config = cfg
                except TypeError:
                    self._ctor_config = {}
            orig(self, *a, **k)

        init._records_config = True
        cls.__init__ = init

    def __init__(self, name: Optional[str] = None, input_shape=None, **_ignored):
        self.name = name
        self.input_shape_arg = tuple(input_shape) if input_shape is not None else None
        self.variables: List[Variable] = []
        self.built = False
        self.in_shape = None   # per-example shape (no batch dim)
        self.out_shape = None
        self.model = None
        self._bufs = {}
        self.need_dx = True    # False for the first weighted layer (no data gradient required)

    # -- construction ------------------------------------------------------------------------
    def add_weight(self, suffix, shape, initializer, trainable=True) -> Variable:
        v = Variable(f"{self.name}/{suffix}", shape, initializer, trainable)
        self.variables.append(v)
        return v

    def build(self, in_shape):  # -> out_shape
        self.in_shape = tuple(in_shape)
        self.out_shape = self.compute_output_shape(self.in_shape)
        self.built = True
        return self.out_shape

    def compute_output_shape(self, in_shape):
        return tuple(in_shape)

    def count_params(self) -> int:
        return sum(v.size for v in self.variables)

    # -- buffers -----------------------------------------------------------------------------
    def buf(self, tag, shape, dtype=torch.float32, device=None):
        key = (tag, tuple(shape), dtype)
        b = self._bufs.get(key)
        if b is None:
            b = torch.empty(tuple(shape), dtype=dtype, device=device if device is not None else self.model.device)
            self._bufs[key] = b
        return b

    def free_buffers(self):
        self._bufs.clear()

    # -- execution ---------------------------------------------------------------------------
    # Activations may be bf16 (policy mixed_bfloat16 with activations='bfloat16', ops/act16.py). ``run`` /
    # ``run_back`` are what executors call: they adapt the dtype at the boundary of a layer that has no
    # bf16 kernel path (a plain copy-cast; the ResNet trunk never takes it), then dispatch.
    def takes_bf16(self, x) -> bool:
        return False

    def run(self, x, training: bool):
        self._in_cast = x.dtype == torch.bfloat16 and not self.takes_bf16(x)
        if self._in_cast:
            x = self.buf("in_f32", x.shape).copy_(x)
        y = self.forward(x, training)
        self._out_dtype = y.dtype
        return y

    def run_back(self, dy):
        want = getattr(self, "_out_dtype", dy.dtype)
        if dy.dtype != want:
            dy = self.buf("dy_cast", dy.shape, want).copy_(dy)
        dx = self.backward(dy)
        if dx is not None and getattr(self, "_in_cast", False):
            dx = self.buf("dx_b16", dx.shape, torch.bfloat16).copy_(dx)
        return dx

    def forward(self, x, training: bool):
        raise NotImplementedError

    def backward(self, dy):
        raise NotImplementedError

    def get_config(self):
        cfg = {}
        for k, v in getattr(self, "_ctor_config", {}).items():
            cfg[k] = list(v) if isinstance(v, tuple) else v
        cfg["name"] = self.name
        return cfg


class InputLayer(Layer):
    _type_name = "input"

    def __init__(self, input_shape, **kw):
        super().__init__(input_shape=input_shape, **kw)

    def forward(self, x, training):
        return x

    def backward(self, dy):
        return dy


class Conv2D(Layer):
    """reference: distributed_with_keras.py:34; mnist_keras_distributed.py:83-98."""
    _type_name = "conv2d"

    def __init__(self, filters, kernel_size, strides=(1, 1), padding="valid", activation=None,
                 use_bias=True, kernel_initializer="glorot_uniform", **kw):
        super().__init__(**kw)
        if activation not in (None, "linear", "relu"):
            raise ValueError(f"Conv2D activation {activation!r} not supported")
        self.filters = int(filters)
        self.kernel_size = _pair(kernel_size)
        self.strides = _pair(strides)
        self.padding = padding.lower()
        self.activation = None if activation == "linear" else activation
        self.use_bias = use_bias

    def compute_output_shape(self, in_shape):
        h, w, _ = in_shape
        ho, wo, self.pads = ops.conv_geometry(h, w, *self.kernel_size, self.strides, self.padding)
        return (ho, wo, self.filters)

    def build(self, in_shape):
        out = super().build(in_shape)
        kh, kw = self.kernel_size
        self.kernel = self.add_weight("kernel", (kh, kw, in_shape[2], self.filters), glorot_uniform)
        self.bias = self.add_weight("bias", (self.filters,), zeros_init) if self.use_bias else None
        return out

    def _act16_ok(self, x) -> bool:
        """bf16-activation path (ops/act16.py): policy mixed_bfloat16 + activations='bfloat16', no fused
        activation, tensor-core-friendly shapes. The fp32 network input is accepted too (stem conv)."""
        from ..ops import mixed, act16
        return (mixed.activation_dtype() == "bfloat16" and self.activation is None
                and act16.conv_supported(x.shape, self.kernel.shape, self.out_shape[0], self.out_shape[1],
                                         x.dtype == torch.bfloat16))

    def takes_bf16(self, x) -> bool:
        return self._act16_ok(x)

    def forward(self, x, training):
        b = x.shape[0]
        self._x = x
        from ..ops import mixed, mixed_conv
        self._a16 = self._act16_ok(x)
        if self._a16:
            from ..ops import act16
            y = self.buf("y16", (b,) + self.out_shape, torch.bfloat16)
            # the gathered im2col operand is kept for the filter-gradient GEMM (HBM is plentiful, bandwidth is not)
            ashape = act16.a_operand_shape(x.shape, self.kernel.shape, self.strides, self.out_shape[0],
                                           self.out_shape[1], x.dtype == torch.bfloat16)
            self._a_cache = (self.buf("a16", ashape, torch.bfloat16)
                             if ashape is not None and training and _CACHE_IM2COL else None)
            # feeding a training-mode batch-norm: the GEMM epilogue leaves its partial statistics behind
            self._bn_partials = None
            if training and x.is_cuda and getattr(self, "feeds_bn", False):
                pshape = act16.bn_partials_shape(x.shape, self.kernel.shape, self.out_shape[0], self.out_shape[1])
                if pshape is not None:
                    self._bn_partials = (self.buf("bn_psum", pshape), self.buf("bn_psq", pshape))
            act16.conv2d_fwd(x, self.kernel.value, self.bias.value if self.bias else None, self.strides,
                             self.pads, out=y, a_cache=self._a_cache, w16=getattr(self.kernel, "value16", None),
                             bn_partials=self._bn_partials)
            self._y = y
            return y
        y = self.buf("y", (b,) + self.out_shape)
        self._mixed = (mixed.global_policy() == "mixed_bfloat16"
                       and mixed_conv.supported(x.shape, self.kernel.shape, self.out_shape[0], self.out_shape[1]))
        if self._mixed:   # tensor-core path: gathered bf16 operands, fp32 accumulation in TMEM
            mixed_conv.conv2d_fwd(x, self.kernel.value, self.bias.value if self.bias else None, self.strides,
                                  self.pads, self.activation == "relu", out=y)
        else:
            ops.conv2d_fwd(x, self.kernel.value, self.bias.value if self.bias else None, self.strides,
                           self.pads, self.activation == "relu", out=y)
        self._y = y
        return y

    def backward(self, dy):
        if getattr(self, "_a16", False):
            from ..ops import act16
            dx = self.buf("dx16", self._x.shape, torch.bfloat16) if self.need_dx else None
            act16.conv2d_bwd(self._x, self.kernel.value, dy, self.strides, self.pads, self.need_dx, self.use_bias,
                             getattr(self, "bias_grad_zero", False), out_dx=dx, out_dw=self.kernel.grad,
                             out_db=self.bias.grad if self.bias else None, a_cache=getattr(self, "_a_cache", None),
                             w16=getattr(self.kernel, "value16", None),
                             side_stream=(self.model.fork_stream() if dy.is_cuda and _FORK_CONV_WGRAD else None))
            if dx is not None and self._x.dtype != torch.bfloat16:
                dx = self.buf("dx", self._x.shape).copy_(dx)
            return dx
        dx = self.buf("dx", self._x.shape) if self.need_dx else None
        if getattr(self, "_mixed", False):
            from ..ops import mixed_conv
            mixed_conv.conv2d_bwd(self._x, self.kernel.value, self._y, dy, self.strides, self.pads,
                                  self.activation == "relu", self.need_dx, self.use_bias, out_dx=dx,
                                  out_dw=self.kernel.grad, out_db=self.bias.grad if self.bias else None)
            return dx
        ops.conv2d_bwd(self._x, self.kernel.value, self._y, dy, self.strides, self.pads,
                       self.activation == "relu", self.need_dx, self.use_bias, out_dx=dx,
                       out_dw=self.kernel.grad, out_db=self.bias.grad if self.bias else None)
        return dx


class MaxPooling2D(Layer):
    """reference: distributed_with_keras.py:35 (defaults: pool 2x2, strides = pool, valid)."""
    _type_name = "max_pooling2d"

    def __init__(self, pool_size=(2, 2), strides=None, padding="valid", **kw):
        super().__init__(**kw)
        self.pool_size = _pair(pool_size)
        self.strides = _pair(strides) if strides is not None else self.pool_size
        self.padding = padding.lower()

    def compute_output_shape(self, in_shape):
        h, w, c = in_shape
        ho, wo, self.pads = ops.conv_geometry(h, w, *self.pool_size, self.strides, self.padding)
        return (ho, wo, c)

    def takes_bf16(self, x) -> bool:
        return x.shape[-1] % 8 == 0

    def forward(self, x, training):
        b = x.shape[0]
        self._x_shape = tuple(x.shape)
        self._b16 = x.dtype == torch.bfloat16
        idx = self.buf("idx", (b,) + self.out_shape, torch.uint8)
        if self._b16:
            from ..ops import act16
            y = self.buf("y16", (b,) + self.out_shape, torch.bfloat16)
            act16.maxpool_fwd(x, self.pool_size, self.strides, self.pads, out=y, out_idx=idx)
        else:
            y = self.buf("y", (b,) + self.out_shape)
            ops.maxpool_fwd(x, self.pool_size, self.strides, self.pads, out=y, out_idx=idx)
        self._idx = idx
        return y

    def backward(self, dy):
        if self._b16:
            from ..ops import act16
            dx = self.buf("dx16", self._x_shape, torch.bfloat16)
            return act16.maxpool_bwd(dy, self._idx, self._x_shape, self.pool_size, self.strides, self.pads, out=dx)
        dx = self.buf("dx", self._x_shape)
        return ops.maxpool_bwd(dy, self._idx, self._x_shape, self.pool_size, self.strides, self.pads, out=dx)


class Flatten(Layer):
    """reference: distributed_with_keras.py:36 -- row-major (h, w, c) order, a free view in NHWC."""
    _type_name = "flatten"

    def takes_bf16(self, x) -> bool:
        return True

    def compute_output_shape(self, in_shape):
        n = 1
        for s in in_shape:
            n *= s
        return (n,)

    def forward(self, x, training):
        self._x_shape = tuple(x.shape)
        return x.reshape(x.shape[0], -1)

    def backward(self, dy):
        return dy.reshape(self._x_shape)


class Reshape(Layer):
    """reference: mnist_keras_distributed.py:81. Accepts any input with the same element count
    ([B,28,28] from the numpy loader or [B,784] from the serving signature)."""
    _type_name = "reshape"

    def __init__(self, target_shape, **kw):
        super().__init__(**kw)
        self.target_shape = tuple(int(s) for s in target_shape)

    def takes_bf16(self, x) -> bool:
        return True

    def compute_output_shape(self, in_shape):
        n = 1
        for s in in_shape:
            n *= s
        t = 1
        for s in self.target_shape:
            t *= s
        if n != t:
            raise ValueError(f"Reshape: cannot reshape {in_shape} into {self.target_shape}")
        return self.target_shape

    def forward(self, x, training):
        self._x_shape = tuple(x.shape)
        return x.reshape((x.shape[0],) + self.target_shape)

    def backward(self, dy):
        return dy.reshape(self._x_shape)


class Dense(Layer):
    """reference: distributed_with_keras.py:37-38; mnist_keras_distributed.py:103,108."""
    _type_name = "dense"

    def __init__(self, units, activation=None, use_bias=True, **kw):
        super().__init__(**kw)
        if activation not in _ACTS:
            raise ValueError(f"Dense activation {activation!r} not supported")
        self.units = int(units)
        self.activation = "none" if activation in (None, "linear") else activation
        self.use_bias = use_bias

    def compute_output_shape(self, in_shape):
        if len(in_shape) != 1:
            raise ValueError("Dense expects a flat input; add Flatten() first")
        return (self.units,)

    def build(self, in_shape):
        out = super().build(in_shape)
        self.kernel = self.add_weight("kernel", (in_shape[0], self.units), glorot_uniform)
        self.bias = self.add_weight("bias", (self.units,), zeros_init) if self.use_bias else None
        return out

    def _mixed_bufs(self, m):
        """bf16 staging buffers of the tensor-core path (static addresses, allocated once per batch size)."""
        k, n = self.kernel.shape
        bf = torch.bfloat16
        return {"xb": self.buf("xb", (m, k), bf), "wt": self.buf("wt", (n, k), bf), "xt": self.buf("xt", (k, m), bf),
                "dyt": self.buf("dyt", (n, m), bf), "dyb": self.buf("dyb", (m, n), bf), "wb": self.buf("wb", (k, n), bf)}

    def forward(self, x, training):
        from ..ops import mixed
        b = x.shape[0]
        self._x = x
        y = self.buf("y", (b, self.units))
        self._mixed = (mixed.global_policy() == "mixed_bfloat16" and self.activation in ("none", "relu")
                       and mixed.supported(b, x.shape[1], self.units))
        if self._mixed:
            # bf16 operands, fp32 accumulation in TMEM (tcgen05); master weights stay fp32
            mixed.dense_fwd(x, self.kernel.value, self.bias.value if self.bias else None, self.activation, out=y,
                            bufs=self._mixed_bufs(b) if x.is_cuda else None)
        else:
            ops.dense_fwd(x, self.kernel.value, self.bias.value if self.bias else None, self.activation, out=y)
        self._y = y
        return y

    def backward(self, dy):
        dx = self.buf("dx", self._x.shape) if self.need_dx else None
        if getattr(self, "_mixed", False):
            from ..ops import mixed
            mixed.dense_bwd(self._x, self.kernel.value, self._y, dy, self.activation, self.need_dx, self.use_bias,
                            out_dx=dx, out_dw=self.kernel.grad, out_db=self.bias.grad if self.bias else None,
                            bufs=self._mixed_bufs(dy.shape[0]) if dy.is_cuda else None)
            return dx
        scratch = self.buf("dz", dy.shape) if self.activation == "softmax" else None
        ops.dense_bwd(self._x, self.kernel.value, self._y, dy, self.activation, self.need_dx, self.use_bias,
                      out_dx=dx, out_dw=self.kernel.grad, out_db=self.bias.grad if self.bias else None,
                      scratch=scratch, side_stream=self.model.fork_stream if dy.is_cuda else None)
        return dx


class BatchNormalization(Layer):
    """reference: mnist_keras_distributed.py:86,93,99,104 (scale=False, center=True). Keras defaults:
    axis=-1, momentum=0.99, epsilon=1e-3. Batch statistics are per replica (no sync-BN), as under
    tf.distribute. ``fused_relu`` is set by the model's fusion pass when an Activation('relu') follows."""
    _type_name = "batch_normalization"

    def __init__(self, axis=-1, momentum=0.99, epsilon=1e-3, center=True, scale=True, **kw):
        super().__init__(**kw)
        if axis not in (-1, 3, 1):
            raise ValueError("BatchNormalization: only the channels-last axis is supported")
        self.momentum, self.epsilon = float(momentum), float(epsilon)
        self.center, self.scale = center, scale
        self.fused_relu = False

    def build(self, in_shape):
        out = super().build(in_shape)
        c = in_shape[-1]
        self.gamma = self.add_weight("gamma", (c,), ones_init) if self.scale else None
        self.beta = self.add_weight("beta", (c,), zeros_init) if self.center else None
        self.moving_mean = self.add_weight("moving_mean", (c,), zeros_init, trainable=False)
        self.moving_variance = self.add_weight("moving_variance", (c,), ones_init, trainable=False)
        # TF's fused kernel (4-D inputs) feeds the unbiased variance to the moving average; the 2-D
        # path feeds the biased one (SURVEY.md section 2.5 parity note).
        self.unbiased_moving = len(in_shape) == 3
        return out

    def takes_bf16(self, x) -> bool:
        from ..ops import act16
        return act16.channels_supported(x.shape[-1])

    def forward(self, x, training):
        c = x.shape[-1]
        self._x = x
        sm = self.buf("save_mean", (c,))
        si = self.buf("save_invstd", (c,))
        self._b16 = x.dtype == torch.bfloat16
        if self._b16:      # bf16 activations: fp32 statistics, folded scale/shift kept for the backward mask
            from ..ops import act16
            y = self.buf("y16", x.shape, torch.bfloat16)
            sc, sh = self.buf("scale", (c,)), self.buf("shift", (c,))
            src = getattr(self, "stats_source", None)      # the convolution whose GEMM already reduced x
            partials = (src._bn_partials if training and src is not None and getattr(src, "_y", None) is x
                        and getattr(src, "_bn_partials", None) is not None else None)
            act16.bn_fwd(x, self.gamma.value if self.gamma else None, self.beta.value if self.beta else None,
                         self.moving_mean.value, self.moving_variance.value, self.momentum, self.epsilon,
                         training, self.fused_relu, self.unbiased_moving, out=y, save_mean=sm, save_invstd=si,
                         scale=sc, shift=sh, partials=partials)
            self._y, self._sm, self._si, self._sc, self._sh = y, sm, si, sc, sh
            return y
        y = self.buf("y", x.shape)
        ops.bn_fwd(x, self.gamma.value if self.gamma else None, self.beta.value if self.beta else None,
                   self.moving_mean.value, self.moving_variance.value, self.momentum, self.epsilon,
                   training, self.fused_relu, self.unbiased_moving, out=y, save_mean=sm, save_invstd=si)
        self._y, self._sm, self._si = y, sm, si
        return y

    def backward(self, dy):
        c = dy.shape[-1]
        dbeta = self.beta.grad if self.beta else self.buf("dbeta", (c,))
        dgamma = self.gamma.grad if self.gamma else self.buf("dgamma", (c,))
        if getattr(self, "_b16", False):
            from ..ops import act16
            dx = self.buf("dx16", self._x.shape, torch.bfloat16)
            act16.bn_bwd(self._x, dy, self.gamma.value if self.gamma else None, self._sm, self._si, self._sc,
                         self._sh, self.fused_relu, out_dx=dx, out_dgamma=dgamma, out_dbeta=dbeta)
            return dx
        dx = self.buf("dx", self._x.shape)
        ops.bn_bwd(self._x, self._y, dy, self.gamma.value if self.gamma else None, self._sm, self._si,
                   self.fused_relu, out_dx=dx, out_dgamma=dgamma, out_dbeta=dbeta)
        return dx


class Activation(Layer):
    """reference: mnist_keras_distributed.py:88. Stand-alone only when it cannot be fused."""
    _type_name = "activation"

    def __init__(self, activation, **kw):
        super().__init__(**kw)
        if activation not in ("relu", "softmax", "linear"):
            raise ValueError(f"Activation {activation!r} not supported")
        self.activation = activation
        self.fused_away = False

    def takes_bf16(self, x) -> bool:
        return self.fused_away or self.activation == "linear"

    def forward(self, x, training):
        if self.fused_away or self.activation == "linear":
            return x
        y = self.buf("y", x.shape)
        if self.activation == "relu":
            ops.relu_fwd(x, out=y)
        else:
            ops.softmax_fwd(x, out=y)
        self._y = y
        return y

    def backward(self, dy):
        if self.fused_away or self.activation == "linear":
            return dy
        dx = self.buf("dx", dy.shape)
        if self.activation == "relu":
            return ops.relu_bwd(self._y, dy, out=dx)
        return ops.softmax_bwd(self._y, dy, out=dx)


class Dropout(Layer):
    """reference: mnist_keras_distributed.py:106. Inverted dropout, identity at inference. The mask is a
    pure function of (seed, step, layer stream id, element) -- nothing is stored for backward."""
    _type_name = "dropout"
    _next_stream = 1

    def __init__(self, rate, seed=None, **kw):
        super().__init__(**kw)
        self.rate = float(rate)
        self.seed = seed
        self.stream_id = 1   # Sequential.add() replaces it with the layer's position in the model

    def forward(self, x, training):
        self._training = training
        if not training or self.rate == 0.0:
            return x
        y = self.buf("y", x.shape)
        m = self.model
        return ops.dropout(x, self.rate, m.seed, m.host_step, self.stream_id + 977 * m.replica_id,
                           hparams=m.hparams, training=True, out=y)

    def backward(self, dy):
        if not self._training or self.rate == 0.0:
            return dy
        dx = self.buf("dx", dy.shape)
        m = self.model
        return ops.dropout(dy, self.rate, m.seed, m.host_step, self.stream_id + 977 * m.replica_id,
                           hparams=m.hparams, training=True, out=dx)


class GlobalAveragePooling2D(Layer):
    _type_name = "global_average_pooling2d"

    def compute_output_shape(self, in_shape):
        return (in_shape[2],)

    def takes_bf16(self, x) -> bool:
        return x.shape[-1] % 8 == 0

    def forward(self, x, training):
        self._x_shape = tuple(x.shape)
        self._b16 = x.dtype == torch.bfloat16
        y = self.buf("y", (x.shape[0], x.shape[3]))
        if self._b16:      # bf16 feature map in, fp32 features out: the classifier head stays fp32
            from ..ops import act16
            return act16.global_avgpool_fwd(x, out=y)
        return ops.global_avgpool_fwd(x, out=y)

    def backward(self, dy):
        if self._b16:
            from ..ops import act16
            return act16.global_avgpool_bwd(dy, self._x_shape, out=self.buf("dx16", self._x_shape, torch.bfloat16))
        return ops.global_avgpool_bwd(dy, self._x_shape, out=self.buf("dx", self._x_shape))


class FusedConvReluPool(Layer):
    """Execution node produced by the fusion pass for Conv2D(3x3, Cin=1, Cout=32, valid, relu) followed
    by MaxPooling2D(2x2): one kernel forward, one kernel backward (csrc/kernels/conv_pool.cu). Owns no
    variables -- it borrows the Conv2D layer's."""
    _type_name = "fused_conv_relu_pool"

    def __init__(self, conv: Conv2D, pool: MaxPooling2D):
        super().__init__(name=f"{conv.name}+{pool.name}")
        self.conv, self.pool = conv, pool
        self.model = conv.model
        self.in_shape, self.out_shape = conv.in_shape, pool.out_shape
        self.in_scale = 1.0 / 255.0

    @staticmethod
    def matches(conv, pool) -> bool:
        return (isinstance(conv, Conv2D) and isinstance(pool, MaxPooling2D)
                and conv.kernel_size == (3, 3) and conv.strides == (1, 1) and conv.padding == "valid"
                and conv.activation == "relu" and conv.use_bias and conv.filters == 32
                and conv.in_shape[2] == 1 and pool.pool_size == (2, 2) and pool.strides == (2, 2)
                and pool.padding == "valid" and not conv.need_dx)

    def forward(self, x, training):
        b = x.shape[0]
        self._x = x
        y = self.buf("y", (b,) + self.out_shape)
        idx = self.buf("idx", (b,) + self.out_shape, torch.uint8)
        ops.conv_relu_pool_fwd(x, self.conv.kernel.value, self.conv.bias.value, self.in_scale, out=y, out_idx=idx)
        self._y, self._idx = y, idx
        return y

    def backward(self, dy):
        ops.conv_relu_pool_bwd(self._x, self.conv.kernel.value, self._y, self._idx, dy, self.in_scale,
                               out_dw=self.conv.kernel.grad, out_db=self.conv.bias.grad)
        return None


def layer_from_config(class_name: str, config: dict) -> Layer:
    """Inverse of ``Layer.get_config`` for the layers a Sequential can hold."""
    table = {c.__name__: c for c in (InputLayer, Conv2D, MaxPooling2D, Flatten, Reshape, Dense,
                                     BatchNormalization, Activation, Dropout, GlobalAveragePooling2D)}
    if class_name not in table:
        raise ValueError(f"unknown layer class {class_name!r}")
    cfg = dict(config)
    for k in ("input_shape", "target_shape", "kernel_size", "strides", "pool_size"):
        if isinstance(cfg.get(k), list):
            cfg[k] = tuple(cfg[k])
    return table[class_name](**cfg)
