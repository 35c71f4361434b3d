"""Convolutions on the tcgen05 bf16 GEMMs (policy ``mixed_bfloat16``).

Each of the three GEMMs of a convolution gets its K-major bf16 operands from a gather kernel
(``csrc/kernels/conv_gather.cu``: im2col, im2col^T, dgrad gather, rotated weights) and runs on
:func:`..tensorcore.gemm_bf16_tn`; outputs, gradients and master weights stay fp32. The gathered operands
live in a grow-only scratch pool shared by all layers (each is consumed by the GEMM right after it is
produced, on one stream), so the extra memory is the largest single operand, not the sum.

CPU restatement (oracle): the fp32 convolution of the bf16-rounded operands.
"""
from __future__ import annotations

import torch

from . import native, tensorcore
from . import reference as ref
from .mixed import colsum_masked, cast_bf16_masked

c_p, c_i, c_l = native.c_p, native.c_i, native.c_l
_G = [c_i] * 13
native.register_optional("tfde_im2col_bf16", [c_p, c_p] + _G + [c_i, c_p])
native.register_optional("tfde_im2colT_bf16", [c_p, c_p] + _G + [c_l, c_p])
native.register_optional("tfde_dgrad_gather_bf16", [c_p, c_p, c_p] + _G + [c_i, c_p])
native.register_optional("tfde_wrot_bf16", [c_p, c_p, c_i, c_i, c_i, c_i, c_i, c_p])
native.register_optional("tfde_cast_transpose_pad_bf16", [c_p, c_p, c_p, c_l, c_i, c_l, c_p])
native.register_optional("tfde_transpose_bf16", [c_p, c_p, c_l, c_i, c_l, c_l, c_p])

_pool = {}


def _scratch(device, tag, numel):
    """Grow-only bf16 scratch, one buffer per (device, stream, tag)."""
    key = (device.index, torch.cuda.current_stream(device).cuda_stream, tag)
    buf = _pool.get(key)
    if buf is None or buf.numel() < numel:
        if torch.cuda.is_current_stream_capturing():
            raise RuntimeError("conv scratch would have to grow during CUDA graph capture; run an eager step first")
        buf = torch.empty(int(numel * 1.0) + 8, dtype=torch.bfloat16, device=device)
        _pool[key] = buf
    return buf[:numel]


def _r8(n):
    return (n + 7) // 8 * 8


def supported(x_shape, w_shape, ho, wo) -> bool:
    b = x_shape[0]
    cout = w_shape[3]
    return (b * ho * wo) % 8 == 0 and cout % 8 == 0 and (b * x_shape[1] * x_shape[2]) % 8 == 0 and b * ho * wo >= 128


def _geom(x_shape, w_shape, stride, pads, ho, wo):
    b, h, wd, c = x_shape
    kh, kw, _, cout = w_shape
    return (b, h, wd, c, kh, kw, ho, wo, stride[0], stride[1], pads[0], pads[2], cout)


def _bf(t):
    return t.to(torch.bfloat16).to(torch.float32)


def conv2d_fwd(x, w, bias=None, stride=(1, 1), pads=(0, 0, 0, 0), relu=False, out=None):
    if not x.is_cuda:
        y = ref.conv2d_fwd(_bf(x), _bf(w), bias, stride, pads, relu)
        if out is not None:
            out.copy_(y)
            return out
        return y
    L = native.load()
    b, h, wd, c = x.shape
    kh, kw, _, cout = w.shape
    ho = (h + pads[0] + pads[1] - kh) // stride[0] + 1
    wo = (wd + pads[2] + pads[3] - kw) // stride[1] + 1
    M, kc = b * ho * wo, kh * kw * c
    kp = _r8(kc)
    dev = x.device
    g = _geom(x.shape, w.shape, stride, pads, ho, wo)
    if kh == 1 and kw == 1 and stride == (1, 1) and kp == kc:
        a = cast_bf16_masked(x.reshape(M, c), None, out=_scratch(dev, "A", M * kp).view(M, kp))
    else:
        a = _scratch(dev, "A", M * kp).view(M, kp)
        native.check(L.tfde_im2col_bf16(x.data_ptr(), a.data_ptr(), *g, kp, native.stream_ptr()), "im2col_bf16")
    wt = _scratch(dev, "B", cout * kp).view(cout, kp)
    native.check(L.tfde_cast_transpose_pad_bf16(w.data_ptr(), None, wt.data_ptr(), kc, cout, kp, native.stream_ptr()),
                 "cast_transpose_pad_bf16")
    y = out if out is not None else torch.empty((b, ho, wo, cout), dtype=torch.float32, device=dev)
    tensorcore.gemm_bf16_tn(a, wt, bias, relu, torch.float32, out=y.view(M, cout))
    return y


def conv2d_bwd(x, w, y, dy, stride=(1, 1), pads=(0, 0, 0, 0), relu=False, need_dx=True, has_bias=True,
               out_dx=None, out_dw=None, out_db=None):
    if not x.is_cuda:
        g = dy * (y > 0).to(dy.dtype) if relu else dy
        dx, dw, db = ref.conv2d_bwd(_bf(x), _bf(w), None, _bf(g), stride, pads, False, need_dx, False)
        db = g.sum(dim=(0, 1, 2)) if has_bias else None
        for o, v in ((out_dx, dx), (out_dw, dw), (out_db, db)):
            if o is not None and v is not None:
                o.copy_(v)
        return (out_dx if out_dx is not None else dx, out_dw if out_dw is not None else dw,
                out_db if out_db is not None else db)
    L = native.load()
    b, h, wd, c = x.shape
    kh, kw, _, cout = w.shape
    ho, wo = dy.shape[1], dy.shape[2]
    M, kc, min_, kd = b * ho * wo, kh * kw * c, b * h * wd, kh * kw * cout
    dev = x.device
    g = _geom(x.shape, w.shape, stride, pads, ho, wo)
    mask = y if relu else None
    dy2 = dy.reshape(M, cout)
    mask2 = mask.reshape(M, cout) if mask is not None else None
    # wgrad: dW[kc, cout] = im2col^T[kc, M] . dy^T[cout, M]^T
    # K-major (M contiguous) im2col operand: gather once in the fast row-major order, then a streaming
    # bf16 transpose (measured ~4x faster than gathering in transposed order)
    kp = _r8(kc)
    if kh == 1 and kw == 1 and stride == (1, 1) and kp == kc:
        a = cast_bf16_masked(x.reshape(M, c), None, out=_scratch(dev, "C", M * kp).view(M, kp))
    else:
        a = _scratch(dev, "C", M * kp).view(M, kp)
        native.check(L.tfde_im2col_bf16(x.data_ptr(), a.data_ptr(), *g, kp, native.stream_ptr()), "im2col_bf16")
    at = _scratch(dev, "A", kc * M).view(kc, M)
    native.check(L.tfde_transpose_bf16(a.data_ptr(), at.data_ptr(), M, kc, kp, M, native.stream_ptr()), "transpose_bf16")
    dyt = _scratch(dev, "B", cout * M).view(cout, M)
    native.check(L.tfde_cast_transpose_pad_bf16(dy2.data_ptr(), native.ptr(mask2), dyt.data_ptr(), M, cout, M,
                                                native.stream_ptr()), "cast_transpose_pad_bf16")
    dw = out_dw if out_dw is not None else torch.empty(tuple(w.shape), dtype=torch.float32, device=dev)
    tensorcore.gemm_bf16_tn(at, dyt, None, False, torch.float32, out=dw.view(kc, cout))
    db = colsum_masked(dy2, mask2, out=out_db) if has_bias else None
    dx = None
    if need_dx:
        if kh == 1 and kw == 1 and stride == (1, 1):
            ga = cast_bf16_masked(dy2, mask2, out=_scratch(dev, "A", M * cout).view(M, cout))     # Min == M
            wr = cast_bf16_masked(w.reshape(c, cout), None, out=_scratch(dev, "B", c * cout).view(c, cout))
        else:
            ga = _scratch(dev, "A", min_ * kd).view(min_, kd)
            native.check(L.tfde_dgrad_gather_bf16(dy.data_ptr(), native.ptr(mask), ga.data_ptr(), *g, kd,
                                                  native.stream_ptr()), "dgrad_gather_bf16")
            wr = _scratch(dev, "B", c * kd).view(c, kd)
            native.check(L.tfde_wrot_bf16(w.data_ptr(), wr.data_ptr(), kh, kw, c, cout, kd, native.stream_ptr()),
                         "wrot_bf16")
        dx = out_dx if out_dx is not None else torch.empty(tuple(x.shape), dtype=torch.float32, device=dev)
        tensorcore.gemm_bf16_tn(ga, wr, None, False, torch.float32, out=dx.view(min_, c))
    return dx, dw, db
