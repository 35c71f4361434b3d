"""Plain-PyTorch restatement of every op the framework runs (the numerical spec).

Used three ways: (1) the CPU execution path (plumbing runs, gloo tests), (2) the oracle the
sm_100a kernels in ``csrc/kernels`` are tested against, (3) documentation of the exact Keras/TF
semantics the reference scripts rely on:

* tensors are NHWC, conv kernels HWIO, dense kernels (in, out) -- the Keras variable layout, which
  is also the [K, N] operand layout of the implicit GEMM (reference: distributed_with_keras.py:33-39,
  mnist_keras_distributed.py:79-109);
* SAME padding follows TensorFlow: ``pad_total = (out-1)*stride + k - in``, low side gets the floor;
* BatchNormalization: eps 1e-3, momentum 0.99, biased batch variance for normalisation;
* SparseCategoricalCrossentropy from logits and from probabilities (clip to [1e-7, 1-1e-7]);
* inverted Dropout driven by a counter-based Philox4x32-10 stream so that the CUDA kernel and this
  file generate the identical mask from (seed, step, stream, element index).

Every function is functional: it returns new tensors and never touches global state.
"""
from __future__ import annotations

import math
from typing import Optional, Sequence, Tuple

import torch
import torch.nn.functional as F

EPS_PROB = 1e-7  # Keras backend epsilon used when the loss is fed probabilities


# ----------------------------------------------------------------------------------------------
# shape helpers
# ----------------------------------------------------------------------------------------------
def same_padding(in_size: int, k: int, stride: int) -> Tuple[int, int, int]:
    """TF 'SAME' rule. Returns (out_size, pad_lo, pad_hi)."""
    out = -(-in_size // stride)
    total = max((out - 1) * stride + k - in_size, 0)
    lo = total // 2
    return out, lo, total - lo


def valid_out(in_size: int, k: int, stride: int) -> int:
    return (in_size - k) // stride + 1


def conv_geometry(h: int, w: int, kh: int, kw: int, stride: Tuple[int, int], padding: str):
    """Returns (Ho, Wo, (pad_top, pad_bottom, pad_left, pad_right)) for 'valid' | 'same'."""
    sh, sw = stride
    if padding == "same":
        ho, pt, pb = same_padding(h, kh, sh)
        wo, pl, pr = same_padding(w, kw, sw)
    elif padding == "valid":
        ho, wo = valid_out(h, kh, sh), valid_out(w, kw, sw)
        pt = pb = pl = pr = 0
    else:
        raise ValueError(f"unknown padding {padding!r}")
    return ho, wo, (pt, pb, pl, pr)


# ----------------------------------------------------------------------------------------------
# Philox4x32-10 (counter based RNG shared with csrc/kernels/elementwise.cu)
# ----------------------------------------------------------------------------------------------
_PHILOX_M0 = 0xD2511F53
_PHILOX_M1 = 0xCD9E8D57
_PHILOX_W0 = 0x9E3779B9
_PHILOX_W1 = 0xBB67AE85
_MASK32 = 0xFFFFFFFF


def philox4x32(c0, c1, c2, c3, k0: int, k1: int):
    """Vectorised Philox4x32-10. Counters are int64 tensors holding uint32 values."""
    for _ in range(10):
        p0 = c0 * _PHILOX_M0  # fits in 64 bit: both < 2^32
        p1 = c2 * _PHILOX_M1
        hi0, lo0 = (p0 >> 32) & _MASK32, p0 & _MASK32
        hi1, lo1 = (p1 >> 32) & _MASK32, p1 & _MASK32
        c0, c1, c2, c3 = (hi1 ^ c1 ^ k0) & _MASK32, lo1, (hi0 ^ c3 ^ k1) & _MASK32, lo0
        k0 = (k0 + _PHILOX_W0) & _MASK32
        k1 = (k1 + _PHILOX_W1) & _MASK32
    return c0, c1, c2, c3


def _mulhilo_safe(a: torch.Tensor, m: int):
    """(a*m) for a < 2^32, m < 2^32 without overflowing signed int64."""
    a_lo, a_hi = a & 0xFFFF, a >> 16
    m_lo, m_hi = m & 0xFFFF, m >> 16
    ll = a_lo * m_lo
    lh = a_lo * m_hi
    hl = a_hi * m_lo
    hh = a_hi * m_hi
    mid = (ll >> 16) + (lh & 0xFFFF) + (hl & 0xFFFF)
    lo = ((mid & 0xFFFF) << 16) | (ll & 0xFFFF)
    hi = hh + (lh >> 16) + (hl >> 16) + (mid >> 16)
    return hi & _MASK32, lo & _MASK32


def philox4x32_safe(c0, c1, c2, c3, k0: int, k1: int):
    for _ in range(10):
        hi0, lo0 = _mulhilo_safe(c0, _PHILOX_M0)
        hi1, lo1 = _mulhilo_safe(c2, _PHILOX_M1)
        c0, c1, c2, c3 = (hi1 ^ c1 ^ k0) & _MASK32, lo1, (hi0 ^ c3 ^ k1) & _MASK32, lo0
        k0 = (k0 + _PHILOX_W0) & _MASK32
        k1 = (k1 + _PHILOX_W1) & _MASK32
    return c0, c1, c2, c3


def dropout_keep_mask(numel: int, rate: float, seed: int, step: int, stream: int,
                      device="cpu") -> torch.Tensor:
    """keep[i] for element i: Philox counter (i//4, step, stream, 0), key (seed_lo, seed_hi),
    lane i%4; keep iff u32 >= rate * 2^32."""
    n4 = (numel + 3) // 4
    idx = torch.arange(n4, dtype=torch.int64, device=device)
    c0 = idx & _MASK32
    c1 = torch.full_like(idx, step & _MASK32)
    c2 = torch.full_like(idx, stream & _MASK32)
    c3 = torch.zeros_like(idx)
    r = philox4x32_safe(c0, c1, c2, c3, seed & _MASK32, (seed >> 32) & _MASK32)
    u = torch.stack(r, dim=1).reshape(-1)[:numel]
    thr = min(int(rate * 4294967296.0), _MASK32)
    return u >= thr


# ----------------------------------------------------------------------------------------------
# convolution (NHWC x HWIO)
# ----------------------------------------------------------------------------------------------
def _nchw(x):
    return x.permute(0, 3, 1, 2)


def _nhwc(x):
    return x.permute(0, 2, 3, 1).contiguous()


def _oihw(w):
    return w.permute(3, 2, 0, 1)


def conv2d_fwd(x, w, bias, stride=(1, 1), pads=(0, 0, 0, 0), relu=False):
    pt, pb, pl, pr = pads
    xp = F.pad(_nchw(x), (pl, pr, pt, pb))
    y = F.conv2d(xp, _oihw(w), bias, stride=stride)
    if relu:
        y = torch.relu(y)
    return _nhwc(y)


def conv2d_bwd(x, w, y, dy, stride=(1, 1), pads=(0, 0, 0, 0), relu=False, need_dx=True,
               has_bias=True):
    """Returns (dx|None, dw, db|None). ``y`` is the saved forward output (ReLU mask)."""
    pt, pb, pl, pr = pads
    if relu:
        dy = dy * (y > 0).to(dy.dtype)
    xp = F.pad(_nchw(x), (pl, pr, pt, pb))
    dyn = _nchw(dy)
    wo = _oihw(w)
    dw = torch.nn.grad.conv2d_weight(xp, wo.shape, dyn, stride=stride)
    dw = dw.permute(2, 3, 1, 0).contiguous()
    db = dy.sum(dim=(0, 1, 2)) if has_bias else None
    dx = None
    if need_dx:
        dxp = torch.nn.grad.conv2d_input(xp.shape, wo, dyn, stride=stride)
        h, wd = x.shape[1], x.shape[2]
        dx = _nhwc(dxp[:, :, pt:pt + h, pl:pl + wd])
    return dx, dw, db


def maxpool_fwd(x, pool=(2, 2), stride=(2, 2), pads=(0, 0, 0, 0)):
    """Returns (y, idx) where idx is the flat index kh*pw+kw of the (first) maximum, uint8."""
    pt, pb, pl, pr = pads
    xn = _nchw(x)
    if any(pads):
        xn = F.pad(xn, (pl, pr, pt, pb), value=float("-inf"))
    n, c, h, w = xn.shape
    ph, pw = pool
    ho, wo = (h - ph) // stride[0] + 1, (w - pw) // stride[1] + 1
    patches = F.unfold(xn, kernel_size=pool, stride=stride)  # [N, C*ph*pw, L]
    patches = patches.reshape(n, c, ph * pw, ho, wo)
    y, idx = patches.max(dim=2)
    return _nhwc(y), idx.permute(0, 2, 3, 1).contiguous().to(torch.uint8)


def maxpool_bwd(dy, idx, x_shape, pool=(2, 2), stride=(2, 2), pads=(0, 0, 0, 0)):
    n, h, w, c = x_shape
    pt, pb, pl, pr = pads
    hp, wp = h + pt + pb, w + pl + pr
    ph, pw = pool
    ho, wo = dy.shape[1], dy.shape[2]
    onehot = F.one_hot(idx.long(), ph * pw).to(dy.dtype)  # [N,Ho,Wo,C,P]
    cols = (onehot * dy.unsqueeze(-1)).permute(0, 3, 4, 1, 2).reshape(n, c * ph * pw, ho * wo)
    dxp = F.fold(cols, (hp, wp), kernel_size=pool, stride=stride)
    return _nhwc(dxp[:, :, pt:pt + h, pl:pl + w])


def conv_relu_pool_fwd(x, w, bias):
    """Fused first layer of Model A: Conv2D(k3, valid, bias, relu) + MaxPooling2D(2x2).
    Reference: distributed_with_keras.py:34-35. Returns (pooled, idx)."""
    y = conv2d_fwd(x, w, bias, relu=True)
    return maxpool_fwd(y)


def conv_relu_pool_bwd(x, w, pooled, idx, dpooled, need_dx=False):
    """Backward of the fused layer: route dpooled to the arg-max position, ReLU mask (pooled>0),
    conv wgrad + bias grad. Returns (dx|None, dw, db)."""
    n, h, wd, _ = x.shape
    kh, kw = w.shape[0], w.shape[1]
    ho, wo = h - kh + 1, wd - kw + 1
    g = dpooled * (pooled > 0).to(dpooled.dtype)
    dy = maxpool_bwd(g, idx, (n, ho, wo, w.shape[3]))
    return conv2d_bwd(x, w, None, dy, relu=False, need_dx=need_dx, has_bias=True)


# ----------------------------------------------------------------------------------------------
# dense
# ----------------------------------------------------------------------------------------------
def softmax_fwd(z):
    return torch.softmax(z, dim=-1)


def softmax_bwd(p, dp):
    return p * (dp - (dp * p).sum(dim=-1, keepdim=True))


def dense_fwd(x, w, bias=None, act: str = "none"):
    y = x @ w
    if bias is not None:
        y = y + bias
    if act == "relu":
        y = torch.relu(y)
    elif act == "softmax":
        y = softmax_fwd(y)
    elif act != "none":
        raise ValueError(act)
    return y


def dense_bwd(x, w, y, dy, act: str = "none", need_dx=True, has_bias=True):
    """Returns (dx|None, dw, db|None)."""
    if act == "relu":
        dy = dy * (y > 0).to(dy.dtype)
    elif act == "softmax":
        dy = softmax_bwd(y, dy)
    dw = x.t() @ dy
    db = dy.sum(dim=0) if has_bias else None
    dx = dy @ w.t() if need_dx else None
    return dx, dw, db


# ----------------------------------------------------------------------------------------------
# batch normalisation (channels last; works for [N,C] and [N,H,W,C])
# ----------------------------------------------------------------------------------------------
def bn_fwd(x, gamma, beta, moving_mean, moving_var, momentum=0.99, eps=1e-3, training=True,
           relu=False, unbiased_moving=True):
    """Returns (y, save_mean, save_invstd, new_moving_mean, new_moving_var).

    ``unbiased_moving``: TF's fused kernel (used for 4-D inputs) feeds the Bessel-corrected batch
    variance into the moving average; the non-fused 2-D path feeds the biased one. The layer picks.
    """
    c = x.shape[-1]
    x2 = x.reshape(-1, c)
    m = x2.shape[0]
    if training:
        mean = x2.mean(dim=0)
        var = x2.var(dim=0, unbiased=False)
        invstd = torch.rsqrt(var + eps)
        var_mov = var * (m / max(m - 1, 1)) if unbiased_moving else var
        new_mm = moving_mean * momentum + mean * (1.0 - momentum)
        new_mv = moving_var * momentum + var_mov * (1.0 - momentum)
    else:
        mean, invstd = moving_mean, torch.rsqrt(moving_var + eps)
        new_mm, new_mv = moving_mean, moving_var
    y = (x2 - mean) * invstd
    if gamma is not None:
        y = y * gamma
    if beta is not None:
        y = y + beta
    if relu:
        y = torch.relu(y)
    return y.reshape(x.shape), mean, invstd, new_mm, new_mv


def bn_bwd(x, y, dy, gamma, save_mean, save_invstd, relu=False):
    """Training-mode backward. Returns (dx, dgamma|None, dbeta)."""
    c = x.shape[-1]
    x2, dy2 = x.reshape(-1, c), dy.reshape(-1, c)
    if relu:
        dy2 = dy2 * (y.reshape(-1, c) > 0).to(dy2.dtype)
    m = x2.shape[0]
    xhat = (x2 - save_mean) * save_invstd
    dbeta = dy2.sum(dim=0)
    dgamma_full = (dy2 * xhat).sum(dim=0)
    g = gamma if gamma is not None else 1.0
    dx = (g * save_invstd / m) * (m * dy2 - dbeta - xhat * dgamma_full)
    return dx.reshape(x.shape), (dgamma_full if gamma is not None else None), dbeta


# ----------------------------------------------------------------------------------------------
# elementwise
# ----------------------------------------------------------------------------------------------
def relu_fwd(x):
    return torch.relu(x)


def relu_bwd(y, dy):
    return dy * (y > 0).to(dy.dtype)


def dropout_fwd(x, rate: float, seed: int, step: int, stream: int):
    keep = dropout_keep_mask(x.numel(), rate, seed, step, stream, device=x.device)
    scale = 1.0 / (1.0 - rate)
    return x * keep.reshape(x.shape).to(x.dtype) * scale


def dropout_bwd(dy, rate: float, seed: int, step: int, stream: int):
    return dropout_fwd(dy, rate, seed, step, stream)


def add(a, b):
    return a + b


def global_avgpool_fwd(x):
    return x.mean(dim=(1, 2))


def global_avgpool_bwd(dy, x_shape):
    n, h, w, c = x_shape
    return (dy / (h * w)).reshape(n, 1, 1, c).expand(n, h, w, c).contiguous()


# ----------------------------------------------------------------------------------------------
# losses (fused with their gradient and the accuracy count)
# ----------------------------------------------------------------------------------------------
def xent_logits(logits, labels, scale: float):
    """SparseCategoricalCrossentropy(from_logits=True) (distributed_with_keras.py:41).
    ``scale`` = 1/global_batch (tf2_mnist_distributed.py:81-83). Returns
    (loss_sum*scale, correct_count, dlogits) with dlogits = (softmax - onehot) * scale."""
    labels = labels.reshape(-1).long()
    bad = (labels < 0) | (labels >= logits.shape[-1])
    if bad.any():
        # a label outside [0, C) poisons the loss with NaN (what TensorFlow's GPU kernel returns) and contributes
        # neither gradient nor accuracy; the CUDA kernels do the same
        safe = labels.masked_fill(bad, 0)
        loss, correct, dlogits = xent_logits(logits[~bad], safe[~bad], scale)
        full = torch.zeros_like(logits)
        full[~bad] = dlogits
        return loss * float("nan"), correct, full
    lse = torch.logsumexp(logits, dim=-1)
    picked = logits.gather(1, labels[:, None]).squeeze(1)
    loss = (lse - picked).sum() * scale
    p = torch.softmax(logits, dim=-1)
    dlogits = (p - F.one_hot(labels, logits.shape[-1]).to(p.dtype)) * scale
    correct = (logits.argmax(dim=-1) == labels).sum().to(logits.dtype)
    return loss, correct, dlogits


def xent_probs(probs, labels, scale: float):
    """'sparse_categorical_crossentropy' on probabilities (mnist_keras_distributed.py:108,114):
    -log(clip(p[y], 1e-7, 1-1e-7)). Returns (loss_sum*scale, correct_count, dprobs)."""
    labels = labels.reshape(-1).long()
    bad = (labels < 0) | (labels >= probs.shape[-1])
    if bad.any():
        safe = labels.masked_fill(bad, 0)
        loss, correct, dp = xent_probs(probs[~bad], safe[~bad], scale)
        full = torch.zeros_like(probs)
        full[~bad] = dp
        return loss * float("nan"), correct, full
    py = probs.gather(1, labels[:, None]).squeeze(1)
    clipped = py.clamp(EPS_PROB, 1.0 - EPS_PROB)
    loss = (-torch.log(clipped)).sum() * scale
    inside = ((py >= EPS_PROB) & (py <= 1.0 - EPS_PROB)).to(probs.dtype)
    dpy = -inside / clipped * scale
    dprobs = torch.zeros_like(probs)
    dprobs.scatter_(1, labels[:, None], dpy[:, None])
    correct = (probs.argmax(dim=-1) == labels).sum().to(probs.dtype)
    return loss, correct, dprobs


# ----------------------------------------------------------------------------------------------
# optimizers on flat arenas (what the fused all-reduce kernel applies as its epilogue)
# ----------------------------------------------------------------------------------------------
def sgd_update(w, g, lr: float, scale: float = 1.0):
    return w - lr * (g * scale)


def momentum_update(w, g, v, lr: float, mu: float, nesterov=False, scale: float = 1.0):
    """Keras SGD(momentum): v = mu*v - lr*g ; w += v  (nesterov: w += mu*v - lr*g)."""
    g = g * scale
    v_new = mu * v - lr * g
    w_new = w + (mu * v_new - lr * g if nesterov else v_new)
    return w_new, v_new


def adam_update(w, g, m, v, lr: float, b1: float, b2: float, eps: float, step: int,
                scale: float = 1.0):
    """Keras Adam: lr_t = lr*sqrt(1-b2^t)/(1-b1^t); w -= lr_t*m/(sqrt(v)+eps). ``step`` is 1-based."""
    g = g * scale
    m_new = b1 * m + (1 - b1) * g
    v_new = b2 * v + (1 - b2) * g * g
    lr_t = lr * math.sqrt(1 - b2 ** step) / (1 - b1 ** step)
    return w - lr_t * m_new / (v_new.sqrt() + eps), m_new, v_new
