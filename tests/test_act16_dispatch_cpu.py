"""The CUDA branch of act16.conv2d_bwd's filter-gradient dispatch, driven on the CPU with the GEMMs / gathers replaced
by PyTorch restatements: which form each layer takes (implicit GEMM, MN-major GEMM on the gathered operand, padded
operand for the 7x7x3 stem, transposing fallback) and that every form produces the same dW as autograd."""
import pytest
import torch
import torch.nn.functional as F

from tfde_b200.ops import act16, conv_igemm, tensorcore

BF = torch.bfloat16


class _AsCuda:
    """A CPU tensor that claims to live on the GPU (only what conv2d_bwd reads before it dispatches)."""

    def __init__(self, t):
        self._t = t

    is_cuda = True

    def __getattr__(self, name):
        return getattr(self._t, name)


def _im2col(x, kh, kw, stride, pads, kp):
    b, h, w, c = x.shape
    xp = F.pad(x.float().permute(0, 3, 1, 2), (pads[2], pads[3], pads[0], pads[1]))
    cols = F.unfold(xp, (kh, kw), stride=stride)                       # [b, c*kh*kw, L] in (c, kh, kw) order
    L = cols.shape[-1]
    cols = cols.view(b, c, kh * kw, L).permute(0, 3, 2, 1).reshape(b * L, kh * kw * c)   # (tap, c) order = HWIO rows
    out = torch.full((b * L, kp), 7.0)                                 # padding columns hold garbage on purpose
    out[:, :kh * kw * c] = cols
    return out.to(BF)


@pytest.mark.parametrize("c,cout,k,nt,expect", [
    (64, 64, 3, True, "igemm"),        # 64-multiple channels: implicit GEMM, nothing gathered
    (16, 32, 3, True, "nt"),           # kc = 144: MN-major GEMM on the gathered operand
    (3, 64, 7, True, "nt_padded"),     # kc = 147 (the stem): padded operand, rows 147..151 dropped
    (16, 32, 3, False, "transpose"),   # TFDE_WGRAD_NT=0: transposing fallback
])
def test_filter_gradient_dispatch(monkeypatch, c, cout, k, nt, expect):
    g = torch.Generator().manual_seed(0)
    b, h = 2, 8
    p = k // 2
    pads, stride = (p, p, p, p), (1, 1)
    x = torch.randn(b, h, h, c, generator=g).to(BF)
    w = torch.randn(k, k, c, cout, generator=g) * 0.1
    dy = torch.randn(b, h, h, cout, generator=g).to(BF)
    taken = []

    def fake_a_operand(x_, w_shape, stride_, pads_, ho, wo, tag, cache=None, reuse=False):
        kp = act16._r8(k * k * c)
        return _im2col(x_._t if isinstance(x_, _AsCuda) else x_, k, k, stride_, pads_, kp), kp

    def fake_nt(a, bmat, out_dtype=torch.float32, out=None):
        taken.append("nt_padded" if a.shape[1] != k * k * c else "nt")
        out.copy_(a.float().t() @ bmat.float())
        return out

    def fake_tn(a, bmat, bias, relu, out_dtype, out=None):
        taken.append("transpose")
        out.copy_(a.float() @ bmat.float().t())
        return out

    def fake_igemm_wgrad(x_, dy_, w_shape, stride_, pads_, out_dw):
        taken.append("igemm")
        kp = k * k * c
        a = _im2col(x_._t, k, k, stride_, pads_, kp)
        out_dw.view(kp, cout).copy_(a.float().t() @ dy_.reshape(-1, cout).float())
        return out_dw

    class _Lib:
        @staticmethod
        def tfde_transpose_bf16(src, dst, rows, cols, ld_src, ld_dst, stream):
            return 0

    def fake_transpose_check(code, what, kernels=1):
        return None

    monkeypatch.setattr(act16, "_a_operand", fake_a_operand)
    monkeypatch.setattr(act16, "_WGRAD_NT", nt)
    monkeypatch.setattr(act16, "_scratch", lambda dev, tag, n: torch.zeros(n, dtype=BF))
    monkeypatch.setattr(tensorcore, "gemm_bf16_nt", fake_nt)
    monkeypatch.setattr(conv_igemm, "wgrad", fake_igemm_wgrad)
    monkeypatch.setattr(conv_igemm, "ENABLED", True)
    monkeypatch.setattr(act16.native, "load", lambda: _Lib)
    monkeypatch.setattr(act16.native, "stream_ptr", lambda: 0)
    if expect == "transpose":
        # the fallback transposes through native kernels: restate them (at = a^T, dyt = dy^T) around the fake GEMM
        kc, M = k * k * c, b * h * h
        a_full = _im2col(x, k, k, stride, pads, act16._r8(kc))

        def fake_tn2(at, dyt, bias, relu, out_dtype, out=None):
            taken.append("transpose")
            out.copy_(a_full[:, :kc].float().t() @ dy.reshape(M, cout).float())
            return out
        monkeypatch.setattr(act16.native, "check", fake_transpose_check)
        monkeypatch.setattr(tensorcore, "gemm_bf16_tn", fake_tn2)
    else:
        monkeypatch.setattr(tensorcore, "gemm_bf16_tn", fake_tn)
    dx, dw, db = act16.conv2d_bwd(_AsCuda(x), w, dy, stride, pads, need_dx=False, has_bias=False)
    assert taken == [expect], taken
    wr = w.to(BF).float().requires_grad_(True)
    xp = F.pad(x.float().permute(0, 3, 1, 2), (p, p, p, p))
    F.conv2d(xp, wr.permute(3, 2, 0, 1)).permute(0, 2, 3, 1).backward(dy.float())
    assert dx is None and db is None
    assert (dw - wr.grad).abs().max().item() <= 1e-3 * wr.grad.abs().max().item()
