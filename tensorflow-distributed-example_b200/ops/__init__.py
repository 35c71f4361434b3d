"""Op layer: one functional API, two executors.

* CUDA tensors  -> hand-written sm_100a kernels (:mod:`.native`, ``csrc/kernels/*.cu``). No fallback:
  a missing/unbuilt library raises.
* CPU tensors   -> :mod:`.reference` (plain PyTorch), which is also the test oracle.

The ``out*`` keyword arguments let the engine hand in preallocated buffers so that a whole training
step has static addresses and can be captured in a CUDA graph.
"""
from __future__ import annotations

import torch

from . import reference as ref
from . import native
from .reference import conv_geometry, same_padding  # noqa: F401


def _cuda(t) -> bool:
    return t.is_cuda


def _put(out, val):
    if out is None:
        return val
    out.copy_(val)
    return out


def conv2d_fwd(x, w, bias=None, stride=(1, 1), pads=(0, 0, 0, 0), relu=False, out=None):
    if _cuda(x):
        return native.conv2d_fwd(x, w, bias, stride, pads, relu, out=out)
    return _put(out, ref.conv2d_fwd(x, w, bias, stride, pads, relu))


def conv2d_bwd(x, w, y, dy, stride=(1, 1), pads=(0, 0, 0, 0), relu=False, need_dx=True, has_bias=True,
               out_dx=None, out_dw=None, out_db=None):
    if _cuda(x):
        return native.conv2d_bwd(x, w, y, dy, stride, pads, relu, need_dx, has_bias, out_dx, out_dw, out_db)
    dx, dw, db = ref.conv2d_bwd(x, w, y, dy, stride, pads, relu, need_dx, has_bias)
    return (_put(out_dx, dx) if dx is not None else None, _put(out_dw, dw),
            _put(out_db, db) if db is not None else None)


def conv_relu_pool_fwd(x, w, bias, in_scale=1.0 / 255.0, out=None, out_idx=None):
    if _cuda(x):
        return native.conv_relu_pool_fwd(x, w, bias, in_scale, out, out_idx)
    xf = x.to(torch.float32) * in_scale if x.dtype == torch.uint8 else x
    p, i = ref.conv_relu_pool_fwd(xf, w, bias)
    return _put(out, p), _put(out_idx, i)


def conv_relu_pool_bwd(x, w, pooled, idx, dpooled, in_scale=1.0 / 255.0, out_dw=None, out_db=None):
    if _cuda(x):
        return native.conv_relu_pool_bwd(x, w, pooled, idx, dpooled, False, in_scale, out_dw, out_db)
    xf = x.to(torch.float32) * in_scale if x.dtype == torch.uint8 else x
    _, dw, db = ref.conv_relu_pool_bwd(xf, w, pooled, idx, dpooled, need_dx=False)
    return None, _put(out_dw, dw), _put(out_db, db)


def maxpool_fwd(x, pool=(2, 2), stride=(2, 2), pads=(0, 0, 0, 0), out=None, out_idx=None):
    if _cuda(x):
        return native.maxpool_fwd(x, pool, stride, pads, out, out_idx)
    y, i = ref.maxpool_fwd(x, pool, stride, pads)
    return _put(out, y), _put(out_idx, i)


def maxpool_bwd(dy, idx, x_shape, pool=(2, 2), stride=(2, 2), pads=(0, 0, 0, 0), out=None):
    if _cuda(dy):
        return native.maxpool_bwd(dy, idx, x_shape, pool, stride, pads, out)
    return _put(out, ref.maxpool_bwd(dy, idx, x_shape, pool, stride, pads))


def dense_fwd(x, w, bias=None, act="none", out=None):
    if _cuda(x):
        return native.dense_fwd(x, w, bias, act, out)
    return _put(out, ref.dense_fwd(x, w, bias, act))


def dense_bwd(x, w, y, dy, act="none", need_dx=True, has_bias=True, out_dx=None, out_dw=None,
              out_db=None, scratch=None, side_stream=None):
    if _cuda(x):
        return native.dense_bwd(x, w, y, dy, act, need_dx, has_bias, out_dx, out_dw, out_db, scratch, side_stream)
    dx, dw, db = ref.dense_bwd(x, w, y, dy, act, need_dx, has_bias)
    return (_put(out_dx, dx) if dx is not None else None, _put(out_dw, dw),
            _put(out_db, db) if db is not None else None)


def bn_fwd(x, gamma, beta, moving_mean, moving_var, momentum=0.99, eps=1e-3, training=True, relu=False,
           unbiased_moving=True, out=None, save_mean=None, save_invstd=None):
    """Updates ``moving_mean`` / ``moving_var`` IN PLACE when training. Returns (y, save_mean, save_invstd)."""
    if _cuda(x):
        y, sm, si, _, _ = native.bn_fwd(x, gamma, beta, moving_mean, moving_var, momentum, eps, training,
                                        relu, unbiased_moving, out, save_mean, save_invstd)
        return y, sm, si
    y, sm, si, mm, mv = ref.bn_fwd(x, gamma, beta, moving_mean, moving_var, momentum, eps, training, relu,
                                   unbiased_moving)
    if training:
        moving_mean.copy_(mm)
        moving_var.copy_(mv)
    return _put(out, y), _put(save_mean, sm), _put(save_invstd, si)


def bn_bwd(x, y, dy, gamma, save_mean, save_invstd, relu=False, out_dx=None, out_dgamma=None,
           out_dbeta=None):
    if _cuda(x):
        return native.bn_bwd(x, y, dy, gamma, save_mean, save_invstd, relu, out_dx, out_dgamma, out_dbeta)
    dx, dg, db = ref.bn_bwd(x, y, dy, gamma, save_mean, save_invstd, relu)
    return _put(out_dx, dx), (_put(out_dgamma, dg) if dg is not None else None), _put(out_dbeta, db)


def relu_fwd(x, out=None):
    return native.relu_fwd(x, out) if _cuda(x) else _put(out, ref.relu_fwd(x))


def relu_bwd(y, dy, out=None):
    return native.relu_bwd(y, dy, out) if _cuda(dy) else _put(out, ref.relu_bwd(y, dy))


def softmax_fwd(z, out=None):
    return native.softmax_fwd(z, out) if _cuda(z) else _put(out, ref.softmax_fwd(z))


def softmax_bwd(p, dp, out=None):
    return native.softmax_bwd(p, dp, out) if _cuda(p) else _put(out, ref.softmax_bwd(p, dp))


def add(a, b, relu=False, out=None):
    if _cuda(a):
        return native.add(a, b, relu, out)
    y = a + b
    return _put(out, torch.relu(y) if relu else y)


def u8_to_f32(x, scale=1.0 / 255.0, out=None):
    if _cuda(x):
        return native.u8_to_f32(x, scale, out)
    return _put(out, x.to(torch.float32) * scale)


def dropout(x, rate, seed, step, stream_id, hparams=None, training=True, out=None):
    """Inverted dropout. On CUDA ``seed``/``step`` are read from the device-side ``hparams`` block (so a
    captured graph advances its own RNG); on CPU they are the explicit arguments."""
    if _cuda(x):
        if hparams is None:
            raise ValueError("CUDA dropout needs the device hparams block")
        return native.dropout(x, rate, hparams, stream_id, training, out)
    if not training:
        return _put(out, x.clone())
    return _put(out, ref.dropout_fwd(x, rate, seed, step, stream_id))


def global_avgpool_fwd(x, out=None):
    return native.global_avgpool_fwd(x, out) if _cuda(x) else _put(out, ref.global_avgpool_fwd(x))


def global_avgpool_bwd(dy, x_shape, out=None):
    return (native.global_avgpool_bwd(dy, x_shape, out) if _cuda(dy)
            else _put(out, ref.global_avgpool_bwd(dy, x_shape)))


def xent(inp, labels, scale, from_probs=False, out_grad=None, out4=None):
    """Returns (out4, dgrad) with out4 = [loss_sum*scale, correct, rows, 0]."""
    if _cuda(inp):
        return native.xent(inp, labels, scale, from_probs, out_grad, out4)
    fn = ref.xent_probs if from_probs else ref.xent_logits
    loss, correct, dg = fn(inp, labels, scale)
    o4 = torch.stack([loss, correct, torch.tensor(float(inp.shape[0]), dtype=inp.dtype),
                      torch.tensor(0.0, dtype=inp.dtype)])
    return _put(out4, o4), _put(out_grad, dg)


def head_supported(h, w) -> bool:
    """True when the fused classifier-head kernel (Dense + loss + its backward in one launch) applies."""
    return _cuda(h) and native.head_supported(int(w.shape[0]), int(w.shape[1]))


def head_fwd_bwd(h, w, bias, labels, scale, from_probs=False, relu_mask=False, need_dh=True, out=None,
                 out_dh=None, out_dw=None, out_db=None, out4=None):
    """Dense(N) -> sparse categorical cross-entropy (+accuracy) -> backward of that Dense.
    ``relu_mask``: the producer of ``h`` had a fused ReLU, so dh is multiplied by (h > 0).
    Returns (out4, out, dh, dw, db); ``out`` holds logits (from_probs=False) or probabilities."""
    if _cuda(h):
        return native.head_fwd_bwd(h, w, bias, labels, scale, from_probs, relu_mask, need_dh, out, out_dh,
                                   out_dw, out_db, out4)
    z = ref.dense_fwd(h, w, bias, "softmax" if from_probs else "none")
    o4, dgrad = xent(z, labels, scale, from_probs)
    dx, dw, db = ref.dense_bwd(h, w, z, dgrad, "softmax" if from_probs else "none", need_dh, bias is not None)
    if dx is not None and relu_mask:
        dx = dx * (h > 0).to(dx.dtype)
    return (_put(out4, o4), _put(out, z) if out is not None else z, _put(out_dh, dx) if dx is not None else None,
            _put(out_dw, dw), _put(out_db, db) if db is not None else None)
