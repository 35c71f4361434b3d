"""Implicit-GEMM convolutions (im2col-mode TMA -> tcgen05) against a plain PyTorch fp32 reference of the same op:
forward, stride-1 data gradient, filter gradient; interior / border / strided / ragged-tile geometries."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(120)]

BF = torch.bfloat16


def _same_pads(n, k, s):
    out = -(-n // s)
    total = max((out - 1) * s + k - n, 0)
    return total // 2, total - total // 2


def _ref(x, w, stride, pads):
    """fp32 convolution of the bf16-rounded operands, NHWC/HWIO in and out (autograd-capable)."""
    xp = F.pad(x.permute(0, 3, 1, 2), (pads[2], pads[3], pads[0], pads[1]))
    return F.conv2d(xp, w.permute(3, 2, 0, 1), stride=stride).permute(0, 2, 3, 1)


# (B, H, W, C, Cout, k, stride, padding)
CASES = [
    (4, 14, 14, 64, 64, 3, 1, "same"),       # ResNet conv5-like tile, M = 784 (ragged last tile)
    (2, 28, 28, 128, 256, 3, 1, "same"),     # two channel blocks per tap, BN = 256
    (3, 7, 7, 64, 128, 3, 1, "same"),        # M = 147: one full and one ragged tile; tensor < 128 KB
    (2, 16, 16, 64, 64, 3, 2, "same"),       # strided, TF-style asymmetric padding (0 before, 1 after)
    (2, 12, 12, 64, 64, 3, 1, "valid"),      # no padding
    (1, 8, 8, 64, 64, 5, 1, "same"),         # 5x5 window, tiny tensor, M = 64 < one tile
    (2, 10, 14, 192, 64, 3, 1, "same"),      # non-square image, three channel blocks
]


def _setup(case, seed=0):
    b, h, wd, c, cout, k, s, padding = case
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(b, h, wd, c, generator=g).to(BF)
    w = (torch.randn(k, k, c, cout, generator=g) * (k * k * c) ** -0.5).to(BF)
    if padding == "same":
        pt, pb = _same_pads(h, k, s)
        pl, pr = _same_pads(wd, k, s)
    else:
        pt = pb = pl = pr = 0
    pads = (pt, pb, pl, pr)
    ho = (h + pt + pb - k) // s + 1
    wo = (wd + pl + pr - k) // s + 1
    dy = torch.randn(b, ho, wo, cout, generator=g).to(BF)
    return x, w, dy, (s, s), pads, ho, wo


@pytest.mark.parametrize("case", CASES)
def test_forward(case):
    from tfde_b200.ops import conv_igemm
    x, w, dy, stride, pads, ho, wo = _setup(case)
    ref = _ref(x.float(), w.float(), stride, pads)
    bias = torch.linspace(-1, 1, w.shape[3])
    y = torch.empty(x.shape[0], ho, wo, w.shape[3], dtype=BF, device="cuda")
    conv_igemm.fwd(x.cuda(), w.cuda(), bias.cuda(), stride, pads, y)
    torch.cuda.synchronize()
    err = (y.float().cpu() - (ref + bias)).abs().max().item()
    scale = (ref + bias).abs().max().item()
    assert err <= 1e-2 * scale, (case, err, scale)


@pytest.mark.parametrize("case", [c for c in CASES if c[6] == 1])
def test_data_gradient(case):
    from tfde_b200.ops import conv_igemm
    x, w, dy, stride, pads, ho, wo = _setup(case, seed=1)
    xr = x.float().requires_grad_(True)
    _ref(xr, w.float(), stride, pads).backward(dy.float())
    dx = torch.empty(tuple(x.shape), dtype=BF, device="cuda")
    conv_igemm.dgrad(dy.cuda(), w.cuda(), pads, dx)
    torch.cuda.synchronize()
    err = (dx.float().cpu() - xr.grad).abs().max().item()
    scale = xr.grad.abs().max().item()
    assert err <= 1e-2 * scale, (case, err, scale)


@pytest.mark.parametrize("case", CASES)
def test_filter_gradient(case):
    from tfde_b200.ops import conv_igemm
    x, w, dy, stride, pads, ho, wo = _setup(case, seed=2)
    wr = w.float().requires_grad_(True)
    _ref(x.float(), wr, stride, pads).backward(dy.float())
    dw = torch.full(tuple(w.shape), float("nan"), dtype=torch.float32, device="cuda")
    conv_igemm.wgrad(x.cuda(), dy.cuda(), tuple(w.shape), stride, pads, dw)
    torch.cuda.synchronize()
    err = (dw.cpu() - wr.grad).abs().max().item()
    scale = wr.grad.abs().max().item()
    assert err <= 2e-4 * scale, (case, err, scale)      # exact bf16 products, fp32 accumulation: order differences only


def test_fused_batchnorm_statistics():
    """The forward GEMM's epilogue leaves per-32-row partial sums / sums of squares of the rounded outputs."""
    from tfde_b200.ops import conv_igemm
    case = (8, 28, 28, 64, 256, 3, 1, "same")
    x, w, dy, stride, pads, ho, wo = _setup(case, seed=3)
    m, cout = x.shape[0] * ho * wo, w.shape[3]
    y = torch.empty(x.shape[0], ho, wo, cout, dtype=BF, device="cuda")
    ps = torch.zeros((m + 31) // 32, cout, device="cuda")
    pq = torch.zeros_like(ps)
    conv_igemm.fwd(x.cuda(), w.cuda(), None, stride, pads, y, (ps, pq))
    torch.cuda.synchronize()
    yf = y.float().view(m, cout)
    assert torch.allclose(ps.sum(0), yf.sum(0), rtol=1e-3, atol=1e-2)
    assert torch.allclose(pq.sum(0), (yf * yf).sum(0), rtol=1e-3, atol=1e-2)
    ref = _ref(x.float(), w.float(), stride, pads)
    assert (yf.cpu().view_as(ref) - ref).abs().max().item() <= 1e-2 * ref.abs().max().item()


def test_layer_path_uses_the_implicit_gemm_and_matches_the_gathered_operand():
    """act16.conv2d_fwd / conv2d_bwd route eligible layers through the implicit GEMM; same numbers as the path that
    materialises the operand (TFDE_CONV_IGEMM=0) up to bf16 rounding of identical fp32 sums."""
    from tfde_b200.ops import act16, conv_igemm, native
    case = (4, 14, 14, 64, 128, 3, 1, "same")
    x, w, dy, stride, pads, ho, wo = _setup(case, seed=4)
    xc, wc, dyc = x.cuda(), w.float().cuda(), dy.cuda()
    names = []
    prev = native.set_trace_sink(names)
    try:
        y1 = act16.conv2d_fwd(xc, wc, None, stride, pads).clone()
        dx1, dw1, _ = act16.conv2d_bwd(xc, wc, dyc, stride, pads, has_bias=False)
        dx1, dw1 = dx1.clone(), dw1.clone()
    finally:
        native.set_trace_sink(prev)
    assert any("conv_igemm_fwd" in str(n) for n in names), names
    assert any("conv_igemm_dgrad" in str(n) for n in names) and any("conv_igemm_wgrad" in str(n) for n in names), names
    old = conv_igemm.ENABLED
    conv_igemm.ENABLED = False
    try:
        y0 = act16.conv2d_fwd(xc, wc, None, stride, pads).clone()
        dx0, dw0, _ = act16.conv2d_bwd(xc, wc, dyc, stride, pads, has_bias=False)
    finally:
        conv_igemm.ENABLED = old
    torch.cuda.synchronize()
    assert (y1.float() - y0.float()).abs().max().item() <= 2e-2 * y0.float().abs().max().item()
    assert (dx1.float() - dx0.float()).abs().max().item() <= 2e-2 * dx0.float().abs().max().item()
    assert (dw1 - dw0).abs().max().item() <= 1e-3 * dw0.abs().max().item()
