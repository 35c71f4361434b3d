"""Implicit-GEMM convolutions on the tcgen05 tensor cores: the im2col operand is never written to HBM, the producer
warp of ``gemm_tcgen05_persistent.cu`` fetches it from the NHWC bf16 activation with im2col-mode TMA loads
(``cuTensorMapEncodeIm2col`` -> ``UTMALDG.4D.IM2COL``).

Forward, data gradient (stride 1) and filter gradient of the reference's Conv2D layers
(/root/reference/mnist_keras_distributed.py:83-98, /root/reference/distributed_with_keras.py:34) for channel counts
that are multiples of 64 (one 128-byte swizzled row per pixel per k-block). Narrower layers keep the direct kernels
(``small_conv.cu``, ``fused_step.cu``) or the gathered operand.

``TFDE_CONV_IGEMM=0`` switches the bf16 convolution path back to the materialised operand (``act16._a_operand``).
"""
from __future__ import annotations

import os

import torch

from . import native

c_p, c_i, c_z = native.c_p, native.c_i, native.c_z
native.register_optional("tfde_conv_igemm_supported", [c_i] * 9)
native.register_optional("tfde_conv_igemm_fwd", [c_p] * 4 + [c_i] * 15 + [c_p, c_p, c_p, c_z, c_p])
native.register_optional("tfde_conv_igemm_dgrad", [c_p] * 3 + [c_i] * 11 + [c_p, c_z, c_p])
native.register_optional("tfde_conv_igemm_wgrad", [c_p] * 3 + [c_i] * 15 + [c_p, c_z, c_p])

ENABLED = os.environ.get("TFDE_CONV_IGEMM", "1") == "1"
BF = torch.bfloat16


def supported(x_shape, w_shape, ho, wo) -> bool:
    """Geometry test only (no GPU needed): 64-multiple channel counts, a real filter window."""
    b, h, wd, c = x_shape
    kh, kw, _, cout = w_shape
    return (kh * kw > 1 and kh <= 16 and kw <= 16 and c % 64 == 0 and cout % 64 == 0 and c > 0 and cout > 0
            and b * ho * wo < 2 ** 31 and b * h * wd < 2 ** 31)


def eligible(x, w_shape, ho, wo) -> bool:
    return bool(ENABLED and x.is_cuda and x.dtype == BF and supported(x.shape, w_shape, ho, wo))


def dgrad_eligible(stride) -> bool:
    return tuple(stride) == (1, 1)


def _w2d(w16, kc, cout):
    if w16.dtype != BF or not w16.is_contiguous():
        raise TypeError("conv_igemm: weights must be a contiguous bf16 HWIO tensor")
    return w16.reshape(kc, cout)


def fwd(x, w16, bias, stride, pads, out, bn_partials=None):
    """out[B,Ho,Wo,Cout] (bf16) = conv(x, w16) + bias; ``bn_partials`` as in act16.conv2d_fwd."""
    L = native.load()
    b, h, wd, c = x.shape
    kh, kw, _, cout = w16.shape
    ho, wo = out.shape[1], out.shape[2]
    if not (x.is_contiguous() and out.is_contiguous()):
        raise ValueError("conv_igemm.fwd: x and out must be contiguous NHWC")
    wsp = native.Workspace.current(x.device)
    ps, pq = (bn_partials[0].data_ptr(), bn_partials[1].data_ptr()) if bn_partials is not None else (None, None)
    native.check(L.tfde_conv_igemm_fwd(x.data_ptr(), _w2d(w16, kh * kw * c, cout).data_ptr(), out.data_ptr(),
                                       native.ptr(bias), b, h, wd, c, kh, kw, ho, wo, stride[0], stride[1],
                                       pads[0], pads[1], pads[2], pads[3], cout, ps, pq, wsp.ws.data_ptr(),
                                       wsp.ws_bytes, native.stream_ptr()), "conv_igemm_fwd")
    return out


def dgrad(dy, w16, pads, out_dx):
    """out_dx[B,H,W,C] (bf16) of a stride-1 convolution with filter w16 [KH,KW,C,Cout] (bf16, as stored)."""
    L = native.load()
    b, h, wd, c = out_dx.shape
    kh, kw, _, cout = w16.shape
    ho, wo = dy.shape[1], dy.shape[2]
    if not (dy.is_contiguous() and out_dx.is_contiguous()):
        raise ValueError("conv_igemm.dgrad: dy and dx must be contiguous NHWC")
    wsp = native.Workspace.current(dy.device)
    native.check(L.tfde_conv_igemm_dgrad(dy.data_ptr(), _w2d(w16, kh * kw * c, cout).data_ptr(), out_dx.data_ptr(),
                                         b, h, wd, c, kh, kw, ho, wo, pads[0], pads[2], cout, wsp.ws.data_ptr(),
                                         wsp.ws_bytes, native.stream_ptr()), "conv_igemm_dgrad")
    return out_dx


def wgrad(x, dy, w_shape, stride, pads, out_dw):
    """out_dw[KH,KW,C,Cout] (fp32) = im2col(x)^T . dy."""
    L = native.load()
    b, h, wd, c = x.shape
    kh, kw, _, cout = w_shape
    ho, wo = dy.shape[1], dy.shape[2]
    if not (x.is_contiguous() and dy.is_contiguous() and out_dw.is_contiguous()):
        raise ValueError("conv_igemm.wgrad: operands must be contiguous")
    wsp = native.Workspace.current(x.device)
    native.check(L.tfde_conv_igemm_wgrad(x.data_ptr(), dy.data_ptr(), out_dw.data_ptr(), b, h, wd, c, kh, kw, ho, wo,
                                         stride[0], stride[1], pads[0], pads[1], pads[2], pads[3], cout,
                                         wsp.ws.data_ptr(), wsp.ws_bytes, native.stream_ptr()), "conv_igemm_wgrad")
    return out_dw
