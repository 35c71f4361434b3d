"""bf16 activations for convolutional networks (``mixed_bfloat16`` with ``activations='bfloat16'``).

Keras' ``mixed_bfloat16`` policy keeps variables in fp32 and runs every layer's compute -- and therefore
every tensor handed from one layer to the next -- in bf16. This module is that activation side:
batch-norm, residual add, ReLU backward, pooling and the convolution GEMM plumbing on bf16 tensors
(``csrc/kernels/act_bf16.cu``), with fp32 statistics / reductions / master weights. For a memory-bound
network this halves (batch-norm backward: x0.38) the bytes every elementwise pass moves and removes the
fp32->bf16 operand casts in front of the tcgen05 GEMMs: a 1x1 convolution reads its A operand straight
from the previous layer's output.

CPU restatement (the oracle for the GPU tests): the same math in fp32 on the bf16-rounded inputs, rounded
to bf16 where the CUDA kernel stores bf16.
"""
from __future__ import annotations

import torch

from . import native, tensorcore
from . import reference as ref
from .mixed import cast_bf16_masked
from .mixed_conv import _scratch, _geom, _r8
from . import conv_igemm

c_p, c_i, c_l, c_f, c_z = native.c_p, native.c_i, native.c_l, native.c_f, native.c_z
BF = torch.bfloat16
_WGRAD_NT = __import__("os").environ.get("TFDE_WGRAD_NT", "1") == "1"   # MN-major filter-gradient GEMM
_FWD_NN = __import__("os").environ.get("TFDE_FWD_NN", "1") == "1"       # forward GEMM on untransposed weights
# BN statistics in the GEMM epilogue: implemented and tested, measured NEUTRAL on ResNet-50 (10.41 vs 10.37 ms/step):
# these GEMMs are output-bound, so the reduction costs in the epilogue what it saves in the statistics kernel (whose
# re-read mostly hits L2). Off by default; TFDE_FUSE_BN_STATS=1 enables it.
_FUSE_BN_STATS = __import__("os").environ.get("TFDE_FUSE_BN_STATS", "0") == "1"

for _n, _a in [
    ("tfde_act16_supported_channels", [c_i]),
    ("tfde_bn_fwd_train_bf16", [c_p, c_p, c_l, c_i, c_p, c_p, c_p, c_p, c_f, c_f, c_i, c_i, c_p, c_p, c_p, c_p,
                                c_p, c_z, c_p, c_p]),
    ("tfde_bn_apply_bf16", [c_p, c_p, c_l, c_i, c_p, c_p, c_i, c_p]),
    ("tfde_bn_fwd_from_partials_bf16", [c_p, c_p, c_l, c_i, c_p, c_p, c_i, c_p, c_p, c_p, c_p, c_f, c_f, c_i, c_i,
                                        c_p, c_p, c_p, c_p, c_p, c_z, c_p, c_p]),
    ("tfde_bn_bwd_bf16", [c_p, c_p, c_p, c_l, c_i, c_p, c_p, c_p, c_p, c_p, c_i, c_p, c_p, c_p, c_z, c_p, c_p]),
    ("tfde_colsum_bf16", [c_p, c_l, c_i, c_p, c_p, c_z, c_p, c_p]),
    ("tfde_add_bf16", [c_p, c_p, c_p, c_l, c_i, c_p]),
    ("tfde_relu_bwd_bf16", [c_p, c_p, c_p, c_p, c_l, c_p]),
    ("tfde_cast_f32_bf16", [c_p, c_p, c_l, c_p]),
    ("tfde_maxpool_fwd_bf16", [c_p, c_p, c_p] + [c_i] * 12 + [c_p]),
    ("tfde_maxpool_bwd_bf16", [c_p, c_p, c_p] + [c_i] * 12 + [c_p]),
    ("tfde_gap_fwd_bf16", [c_p, c_p, c_i, c_i, c_i, c_p]),
    ("tfde_gap_bwd_bf16", [c_p, c_p, c_i, c_i, c_i, c_p]),
    ("tfde_im2col_b2b", [c_p, c_p] + [c_i] * 13 + [c_i, c_p]),
    ("tfde_dgrad_gather_b2b", [c_p, c_p] + [c_i] * 13 + [c_i, c_p]),
    ("tfde_strided_scatter_b2b", [c_p, c_p] + [c_i] * 8 + [c_p]),
]:
    native.register_optional(_n, _a)


def channels_supported(c: int) -> bool:
    """Channel counts the 8-wide kernels take: C % 8 == 0 and (C/8 divides 256 or C % 256 == 0)."""
    if c < 8 or c % 8:
        return False
    t = min(c // 8, 32)
    return 256 % t == 0 and c % (8 * t) == 0


def _ws(device):
    # partials of the statistics kernels live in the big per-stream scratch (shared, in stream order, with the
    # split-K GEMMs): 2048 channels x 111 row blocks x 2 sums do not fit the 1 MB `small` buffer
    w = native.Workspace.current(device)
    return w, w.ws.data_ptr(), w.ws.numel(), w.counters.data_ptr()


def _empty_like_bf(t, out):
    return out if out is not None else torch.empty(t.shape, dtype=BF, device=t.device)


# ----------------------------------------------------------------------------------------------
# batch normalisation
# ----------------------------------------------------------------------------------------------
def bn_fwd(x, gamma, beta, moving_mean, moving_var, momentum, eps, training, relu, unbiased_moving,
           out=None, save_mean=None, save_invstd=None, scale=None, shift=None, partials=None):
    """x, out: bf16 [..., C]. Updates the moving statistics in place when training. ``partials`` = (psum, psq)
    per-32-row partial statistics left by the producing GEMM (CUDA): x is then read once. Returns
    (y, save_mean, save_invstd, scale, shift); scale/shift are the folded per-channel affine that the
    backward uses to recompute the fused-ReLU mask from x."""
    C = x.shape[-1]
    M = x.numel() // C
    dev = x.device
    f = lambda t: t if t is not None else torch.empty(C, dtype=torch.float32, device=dev)
    save_mean, save_invstd, scale, shift = f(save_mean), f(save_invstd), f(scale), f(shift)
    if not x.is_cuda:
        xf = x.float()
        if training:
            y, sm, si, mm, mv = ref.bn_fwd(xf, gamma, beta, moving_mean, moving_var, momentum, eps, True, False,
                                           unbiased_moving)
            moving_mean.copy_(mm); moving_var.copy_(mv)
            save_mean.copy_(sm); save_invstd.copy_(si)
        else:
            save_mean.copy_(moving_mean); save_invstd.copy_(torch.rsqrt(moving_var + eps))
        sc = save_invstd * (gamma if gamma is not None else 1.0)
        scale.copy_(sc)
        shift.copy_((beta if beta is not None else 0.0) - save_mean * sc)
        y = torch.addcmul(shift, xf, scale)
        y = torch.relu(y) if relu else y
        y = y.to(BF)
        if out is not None:
            out.copy_(y)
            y = out
        return y, save_mean, save_invstd, scale, shift
    L = native.load()
    y = _empty_like_bf(x, out)
    if training and partials is not None:
        psum, psq = partials
        _, ws, wsn, cnt = _ws(dev)
        native.check(L.tfde_bn_fwd_from_partials_bf16(x.data_ptr(), y.data_ptr(), M, C, psum.data_ptr(), psq.data_ptr(),
                                                      psum.shape[0], native.ptr(gamma), native.ptr(beta),
                                                      moving_mean.data_ptr(), moving_var.data_ptr(), momentum, eps,
                                                      1 if relu else 0, 1 if unbiased_moving else 0,
                                                      save_mean.data_ptr(), save_invstd.data_ptr(), scale.data_ptr(),
                                                      shift.data_ptr(), ws, wsn, cnt, native.stream_ptr()),
                     "bn_fwd_from_partials_bf16")
    elif training:
        _, ws, wsn, cnt = _ws(dev)
        native.check(L.tfde_bn_fwd_train_bf16(x.data_ptr(), y.data_ptr(), M, C, native.ptr(gamma), native.ptr(beta),
                                              moving_mean.data_ptr(), moving_var.data_ptr(), momentum, eps,
                                              1 if relu else 0, 1 if unbiased_moving else 0, save_mean.data_ptr(),
                                              save_invstd.data_ptr(), scale.data_ptr(), shift.data_ptr(), ws, wsn,
                                              cnt, native.stream_ptr()), "bn_fwd_train_bf16")
    else:
        save_mean.copy_(moving_mean)
        torch.rsqrt(moving_var + eps, out=save_invstd)
        torch.mul(save_invstd, gamma, out=scale) if gamma is not None else scale.copy_(save_invstd)
        shift.copy_((beta if beta is not None else 0.0) - save_mean * scale)
        native.check(L.tfde_bn_apply_bf16(x.data_ptr(), y.data_ptr(), M, C, scale.data_ptr(), shift.data_ptr(),
                                          1 if relu else 0, native.stream_ptr()), "bn_apply_bf16")
    return y, save_mean, save_invstd, scale, shift


def bn_bwd(x, dy, gamma, save_mean, save_invstd, scale, shift, relu, out_dx=None, out_dgamma=None, out_dbeta=None):
    """Returns (dx bf16, dgamma, dbeta). With ``relu`` the mask (y > 0) is recomputed as x*scale+shift > 0."""
    C = x.shape[-1]
    M = x.numel() // C
    dev = x.device
    dgamma = out_dgamma if out_dgamma is not None else torch.empty(C, dtype=torch.float32, device=dev)
    dbeta = out_dbeta if out_dbeta is not None else torch.empty(C, dtype=torch.float32, device=dev)
    if not x.is_cuda:
        xf, g = x.float(), dy.float()
        if relu:
            g = g * (torch.addcmul(shift, xf, scale) > 0).to(g.dtype)
        dx, dg, db = ref.bn_bwd(xf, None, g, gamma, save_mean, save_invstd, False)
        dgamma.copy_(dg if dg is not None else (g * (xf - save_mean) * save_invstd).reshape(-1, C).sum(0))
        dbeta.copy_(db)
        dx = dx.to(BF)
        if out_dx is not None:
            out_dx.copy_(dx)
            dx = out_dx
        return dx, dgamma, dbeta
    L = native.load()
    dx = _empty_like_bf(x, out_dx)
    _, ws, wsn, cnt = _ws(dev)
    native.check(L.tfde_bn_bwd_bf16(x.data_ptr(), dy.data_ptr(), dx.data_ptr(), M, C, native.ptr(gamma),
                                    save_mean.data_ptr(), save_invstd.data_ptr(), scale.data_ptr(), shift.data_ptr(),
                                    1 if relu else 0, dgamma.data_ptr(), dbeta.data_ptr(), ws, wsn, cnt,
                                    native.stream_ptr()), "bn_bwd_bf16")
    return dx, dgamma, dbeta


def colsum(x2, out=None):
    """fp32 column sums of a bf16 [M, C] matrix (bias gradient)."""
    M, C = x2.shape
    o = out if out is not None else torch.empty(C, dtype=torch.float32, device=x2.device)
    if not x2.is_cuda:
        o.copy_(x2.float().sum(0))
        return o
    _, ws, wsn, cnt = _ws(x2.device)
    native.check(native.load().tfde_colsum_bf16(x2.data_ptr(), M, C, o.data_ptr(), ws, wsn, cnt, native.stream_ptr()),
                 "colsum_bf16")
    return o


# ----------------------------------------------------------------------------------------------
# elementwise / pooling
# ----------------------------------------------------------------------------------------------
def add(a, b, relu=False, out=None):
    if not a.is_cuda:
        y = a.float() + b.float()
        y = (torch.relu(y) if relu else y).to(BF)
        if out is not None:
            out.copy_(y)
            return out
        return y
    y = _empty_like_bf(a, out)
    native.check(native.load().tfde_add_bf16(a.data_ptr(), b.data_ptr(), y.data_ptr(), a.numel(), 1 if relu else 0,
                                             native.stream_ptr()), "add_bf16")
    return y


def relu_bwd(y, dy, addend=None, out=None):
    """dx = dy * (y > 0) [+ addend]."""
    if not y.is_cuda:
        d = dy.float() * (y.float() > 0).float()
        if addend is not None:
            d = d + addend.float()
        d = d.to(BF)
        if out is not None:
            out.copy_(d)
            return out
        return d
    dx = _empty_like_bf(y, out)
    native.check(native.load().tfde_relu_bwd_bf16(y.data_ptr(), dy.data_ptr(), native.ptr(addend), dx.data_ptr(),
                                                  y.numel(), native.stream_ptr()), "relu_bwd_bf16")
    return dx


def cast_to_bf16(x, out=None):
    if not x.is_cuda:
        y = x.to(BF)
        if out is not None:
            out.copy_(y)
            return out
        return y
    y = _empty_like_bf(x, out)
    native.check(native.load().tfde_cast_f32_bf16(x.data_ptr(), y.data_ptr(), x.numel(), native.stream_ptr()),
                 "cast_f32_bf16")
    return y


def maxpool_fwd(x, pool, stride, pads, out=None, out_idx=None):
    b, h, w, c = x.shape
    ho = (h + pads[0] + pads[1] - pool[0]) // stride[0] + 1
    wo = (w + pads[2] + pads[3] - pool[1]) // stride[1] + 1
    if not x.is_cuda:
        y, i = ref.maxpool_fwd(x.float(), pool, stride, pads)
        y = y.to(BF)
        if out is not None:
            out.copy_(y); y = out
        if out_idx is not None:
            out_idx.copy_(i); i = out_idx
        return y, i
    y = out if out is not None else torch.empty((b, ho, wo, c), dtype=BF, device=x.device)
    idx = out_idx if out_idx is not None else torch.empty((b, ho, wo, c), dtype=torch.uint8, device=x.device)
    native.check(native.load().tfde_maxpool_fwd_bf16(x.data_ptr(), y.data_ptr(), idx.data_ptr(), b, h, w, c, pool[0],
                                                     pool[1], stride[0], stride[1], pads[0], pads[2], ho, wo,
                                                     native.stream_ptr()), "maxpool_fwd_bf16")
    return y, idx


def maxpool_bwd(dy, idx, x_shape, pool, stride, pads, out=None):
    b, h, w, c = x_shape
    ho, wo = dy.shape[1], dy.shape[2]
    if not dy.is_cuda:
        dx = ref.maxpool_bwd(dy.float(), idx, x_shape, pool, stride, pads).to(BF)
        if out is not None:
            out.copy_(dx)
            return out
        return dx
    dx = out if out is not None else torch.empty(tuple(x_shape), dtype=BF, device=dy.device)
    native.check(native.load().tfde_maxpool_bwd_bf16(dy.data_ptr(), idx.data_ptr(), dx.data_ptr(), b, h, w, c, pool[0],
                                                     pool[1], stride[0], stride[1], pads[0], pads[2], ho, wo,
                                                     native.stream_ptr()), "maxpool_bwd_bf16")
    return dx


def global_avgpool_fwd(x, out=None):
    """bf16 [B, H, W, C] -> fp32 [B, C] (the classifier head stays fp32)."""
    b, h, w, c = x.shape
    if not x.is_cuda:
        y = x.float().mean(dim=(1, 2))
        if out is not None:
            out.copy_(y)
            return out
        return y
    y = out if out is not None else torch.empty((b, c), dtype=torch.float32, device=x.device)
    native.check(native.load().tfde_gap_fwd_bf16(x.data_ptr(), y.data_ptr(), b, h * w, c, native.stream_ptr()),
                 "gap_fwd_bf16")
    return y


def global_avgpool_bwd(dy, x_shape, out=None):
    b, h, w, c = x_shape
    if not dy.is_cuda:
        dx = (dy.float() / float(h * w)).reshape(b, 1, 1, c).expand(b, h, w, c).to(BF).contiguous()
        if out is not None:
            out.copy_(dx)
            return out
        return dx
    dx = out if out is not None else torch.empty(tuple(x_shape), dtype=BF, device=dy.device)
    native.check(native.load().tfde_gap_bwd_bf16(dy.data_ptr(), dx.data_ptr(), b, h * w, c, native.stream_ptr()),
                 "gap_bwd_bf16")
    return dx


# ----------------------------------------------------------------------------------------------
# convolution: bf16 activations in, bf16 activations out, fp32 weights / weight gradients
# ----------------------------------------------------------------------------------------------
def conv_supported(x_shape, w_shape, ho, wo, x_is_bf16=True) -> bool:
    b, h, wd, c = x_shape
    cout = w_shape[3]
    M, Min = b * ho * wo, b * h * wd
    return (M % 8 == 0 and Min % 8 == 0 and M >= 128 and channels_supported(cout)
            and (not x_is_bf16 or c % 8 == 0))


def _bf(t):
    return t.to(BF).to(torch.float32)


def a_operand_shape(x_shape, w_shape, stride, ho, wo, x_is_bf16=True):
    """(M, kp) of the gathered A operand, or None when the activation itself is the operand (1x1, stride 1)."""
    b, h, wd, c = x_shape
    kh, kw = w_shape[0], w_shape[1]
    if x_is_bf16 and kh == 1 and kw == 1 and tuple(stride) == (1, 1):
        return None
    if x_is_bf16 and conv_igemm.ENABLED and conv_igemm.supported(x_shape, w_shape, ho, wo):
        return None                                    # implicit GEMM: the operand is fetched by im2col-mode TMA
    return (b * ho * wo, _r8(kh * kw * c))


def _a_operand(x, w_shape, stride, pads, ho, wo, tag, cache=None, reuse=False):
    """K-major bf16 A operand [M, kp] of the forward / filter-gradient GEMM. ``cache``: a caller-owned
    [M, kp] bf16 buffer to gather into (forward) or, with ``reuse``, to take as is (backward: the im2col
    matrix of the forward pass is kept instead of being gathered a second time -- memory for time)."""
    L = native.load()
    b, h, wd, c = x.shape
    kh, kw, _, cout = w_shape
    M, kc = b * ho * wo, kh * kw * c
    kp = _r8(kc)
    g = _geom(x.shape, w_shape, stride, pads, ho, wo)
    if x.dtype == BF:
        if kh == 1 and kw == 1 and tuple(stride) == (1, 1):
            return x.reshape(M, c), kp                 # the activation IS the operand
        if cache is not None and reuse:
            return cache, kp
        a = cache if cache is not None else _scratch(x.device, tag, M * kp).view(M, kp)
        native.check(L.tfde_im2col_b2b(x.data_ptr(), a.data_ptr(), *g, kp, native.stream_ptr()), "im2col_b2b")
        return a, kp
    if cache is not None and reuse:
        return cache, kp
    a = cache if cache is not None else _scratch(x.device, tag, M * kp).view(M, kp)
    native.check(L.tfde_im2col_bf16(x.data_ptr(), a.data_ptr(), *g, kp, native.stream_ptr()), "im2col_bf16")
    return a, kp


def bn_partials_shape(x_shape, w_shape, ho, wo):
    """[G, cout] of the partial-statistics buffers conv2d_fwd can fill for a following batch-norm, or None when
    the fused-statistics GEMM does not apply (small outputs keep split-K and the separate statistics kernel)."""
    M, cout = x_shape[0] * ho * wo, w_shape[3]
    if not (_FWD_NN and _FUSE_BN_STATS and tensorcore.nn_stats_eligible(M, cout)):
        return None
    return ((M + 31) // 32, cout)


def conv2d_fwd(x, w, bias=None, stride=(1, 1), pads=(0, 0, 0, 0), out=None, a_cache=None, w16=None,
               bn_partials=None):
    """x: bf16 NHWC (or the fp32 network input); w: fp32 HWIO; returns bf16 NHWC.
    ``bn_partials`` = (psum, psq) fp32 [G, cout] buffers (see bn_partials_shape): CUDA only, filled by the GEMM."""
    b, h, wd, c = x.shape
    kh, kw, _, cout = w.shape
    ho = (h + pads[0] + pads[1] - kh) // stride[0] + 1
    wo = (wd + pads[2] + pads[3] - kw) // stride[1] + 1
    if not x.is_cuda:
        y = ref.conv2d_fwd(_bf(x.float()), _bf(w), bias, stride, pads, False).to(BF)
        if out is not None:
            out.copy_(y)
            return out
        return y
    L = native.load()
    M, kc = b * ho * wo, kh * kw * c
    y = out if out is not None else torch.empty((b, ho, wo, cout), dtype=BF, device=x.device)
    if conv_igemm.eligible(x, w.shape, ho, wo):
        # implicit GEMM: A tiles come straight from the activation (im2col-mode TMA), B is the bf16 weight copy
        wb = w16 if w16 is not None else cast_bf16_masked(w.reshape(kc, cout), None,
                                                          out=_scratch(x.device, "B", kc * cout).view(kc, cout))
        return conv_igemm.fwd(x, wb.reshape(kh, kw, c, cout), bias, stride, pads, y, bn_partials)
    a, kp = _a_operand(x, w.shape, stride, pads, ho, wo, "A", cache=a_cache)
    if _FWD_NN:
        # weights exactly as Keras stores them (HWIO = [kc, cout] row-major): the GEMM reads them MN-major, the
        # only per-step weight work is a flat fp32 -> bf16 cast (padded im2col columns meet zero-filled rows)
        wb = w16 if w16 is not None else cast_bf16_masked(w.reshape(kc, cout), None,
                                                          out=_scratch(x.device, "B", kc * cout).view(kc, cout))
        if bn_partials is not None:
            # the consumer is a training-mode batch-norm: leave its per-32-row partial statistics behind
            tensorcore.gemm_bf16_nn_stats(a, wb.reshape(kc, cout), bias, y.view(M, cout), bn_partials[0], bn_partials[1])
        else:
            tensorcore.gemm_bf16_nn(a, wb.reshape(kc, cout), bias, False, BF, out=y.view(M, cout))
        return y
    wt = _scratch(x.device, "B", cout * kp).view(cout, kp)
    native.check(L.tfde_cast_transpose_pad_bf16(w.data_ptr(), None, wt.data_ptr(), kc, cout, kp, native.stream_ptr()),
                 "cast_transpose_pad_bf16")
    tensorcore.gemm_bf16_tn(a, wt, bias, False, BF, out=y.view(M, cout))
    return y


def conv2d_bwd(x, w, dy, stride=(1, 1), pads=(0, 0, 0, 0), need_dx=True, has_bias=True, bias_grad_zero=False,
               out_dx=None, out_dw=None, out_db=None, a_cache=None, side_stream=None, w16=None):
    """dy: bf16. Returns (dx bf16 | None, dw fp32, db fp32 | None).
    ``bias_grad_zero``: the convolution feeds a training-mode batch-norm, whose input gradient sums to zero
    over every channel (batch-norm is invariant to a per-channel shift of its input), so the bias gradient
    is identically zero and its reduction is skipped."""
    b, h, wd, c = x.shape
    kh, kw, _, cout = w.shape
    ho, wo = dy.shape[1], dy.shape[2]
    if not x.is_cuda:
        g = dy.float()
        dx, dw, _ = ref.conv2d_bwd(_bf(x.float()), _bf(w), None, g, stride, pads, False, need_dx, False)
        db = None
        if has_bias:
            db = torch.zeros(cout) if bias_grad_zero else g.sum(dim=(0, 1, 2))
        if dx is not None:
            dx = dx.to(BF)
        for o, v in ((out_dx, dx), (out_dw, dw), (out_db, db)):
            if o is not None and v is not None:
                o.copy_(v)
        return (out_dx if out_dx is not None and dx is not None else dx, out_dw if out_dw is not None else dw,
                out_db if out_db is not None and db is not None else db)
    L = native.load()
    dev = x.device
    M, kc, min_, kd = b * ho * wo, kh * kw * c, b * h * wd, kh * kw * cout
    dy2 = dy.reshape(M, cout)
    # filter gradient: dW[kc, cout] = A^T[kc, M] . dy^T[cout, M]^T (both K-major after a bf16 transpose)
    dw = out_dw if out_dw is not None else torch.empty(tuple(w.shape), dtype=torch.float32, device=dev)
    igemm = conv_igemm.eligible(x, w.shape, ho, wo) and dy.dtype == BF and dy.is_contiguous()
    if igemm:
        # implicit GEMM: im2col(x) tiles fetched MN-major by TMA, dy as it is, reduced over the output pixels
        if side_stream is not None and need_dx:
            side_stream.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side_stream):
                conv_igemm.wgrad(x, dy, w.shape, stride, pads, dw)
        else:
            conv_igemm.wgrad(x, dy, w.shape, stride, pads, dw)
    else:
        a, kp = _a_operand(x, w.shape, stride, pads, ho, wo, "C", cache=a_cache, reuse=a_cache is not None)
        if _WGRAD_NT and kc % 8 == 0:
            # operands as the forward pass left them: A[M, kc] and dy[M, cout], reduced over rows -> the tcgen05
            # GEMM reads them MN-major, no transposed copies
            a_nt = a[:, :kc] if kp != kc else a
            owns_a = a_cache is not None or a.data_ptr() == x.data_ptr()     # not in the shared scratch pool
            if side_stream is not None and need_dx and owns_a:
                # only the optimizer consumes dW: fork it so it overlaps the data-gradient chain of this layer and
                # the batch-norm backward of the previous one (joined before the optimizer / all-reduce)
                side_stream.wait_stream(torch.cuda.current_stream(dev))
                with torch.cuda.stream(side_stream):
                    tensorcore.gemm_bf16_nt(a_nt, dy2, out=dw.view(kc, cout))
            else:
                tensorcore.gemm_bf16_nt(a_nt, dy2, out=dw.view(kc, cout))
        elif _WGRAD_NT and cout % 8 == 0:
            # kc not a multiple of 8 (the 7x7x3 stem): reduce the PADDED operand [M, kp] the same way -- rows kc..kp of
            # the product come from the padding columns and are dropped -- instead of transposing both operands
            dwp = _scratch(dev, "Wp", kp * cout * 2).view(torch.float32).view(kp, cout)
            tensorcore.gemm_bf16_nt(a, dy2, out=dwp)
            dw.view(kc, cout).copy_(dwp[:kc])
        else:
            at = _scratch(dev, "A", kc * M).view(kc, M)
            native.check(L.tfde_transpose_bf16(a.data_ptr(), at.data_ptr(), M, kc, kp, M, native.stream_ptr()),
                         "transpose_bf16")
            dyt = _scratch(dev, "B", cout * M).view(cout, M)
            native.check(L.tfde_transpose_bf16(dy2.data_ptr(), dyt.data_ptr(), M, cout, cout, M, native.stream_ptr()),
                         "transpose_bf16")
            tensorcore.gemm_bf16_tn(at, dyt, None, False, torch.float32, out=dw.view(kc, cout))
    db = None
    if has_bias:
        if bias_grad_zero:
            db = out_db if out_db is not None else torch.zeros(cout, dtype=torch.float32, device=dev)
        else:
            db = colsum(dy2, out=out_db)
    dx = None
    if need_dx:
        dx = out_dx if out_dx is not None else torch.empty(tuple(x.shape), dtype=BF, device=dev)
        if kh == 1 and kw == 1 and tuple(stride) != (1, 1) and tuple(pads) == (0, 0, 0, 0) and M >= 128:
            # strided 1x1 (projection shortcuts): GEMM on the compact [M, c] problem, then spread the rows over the
            # input grid with zeros in between -- not a [B*H*W, cout] operand that is 3/4 zeros
            wr = (w16.reshape(c, cout) if w16 is not None else
                  cast_bf16_masked(w.reshape(c, cout), None, out=_scratch(dev, "B", c * cout).view(c, cout)))
            gc = _scratch(dev, "A", M * c).view(M, c)
            tensorcore.gemm_bf16_tn(dy2, wr, None, False, BF, out=gc)
            native.check(L.tfde_strided_scatter_b2b(gc.data_ptr(), dx.data_ptr(), b, h, wd, c, ho, wo, stride[0],
                                                    stride[1], native.stream_ptr()), "strided_scatter_b2b")
            return dx, dw, db
        if igemm and conv_igemm.dgrad_eligible(stride):
            # a stride-1 data gradient is a convolution of dy with the mirrored filter: same kernel, the forward
            # pass's bf16 weights addressed tap-mirrored (no rotated copy, no gathered operand)
            wb = w16 if w16 is not None else cast_bf16_masked(w.reshape(kc, cout), None,
                                                              out=_scratch(dev, "B", kc * cout).view(kc, cout))
            conv_igemm.dgrad(dy, wb.reshape(kh, kw, c, cout), pads, dx)
            return dx, dw, db
        if kh == 1 and kw == 1 and tuple(stride) == (1, 1):
            ga = dy2                                                     # Min == M: dy IS the operand
            wr = (w16.reshape(c, cout) if w16 is not None else
                  cast_bf16_masked(w.reshape(c, cout), None, out=_scratch(dev, "B", c * cout).view(c, cout)))
        else:
            ga = _scratch(dev, "A", min_ * kd).view(min_, kd)
            g = _geom(x.shape, w.shape, stride, pads, ho, wo)
            native.check(L.tfde_dgrad_gather_b2b(dy.data_ptr(), ga.data_ptr(), *g, kd, native.stream_ptr()),
                         "dgrad_gather_b2b")
            wr = _scratch(dev, "B", c * kd).view(c, kd)
            native.check(L.tfde_wrot_bf16(w.data_ptr(), wr.data_ptr(), kh, kw, c, cout, kd, native.stream_ptr()),
                         "wrot_bf16")
        tensorcore.gemm_bf16_tn(ga, wr, None, False, BF, out=dx.view(min_, c))
    return dx, dw, db
