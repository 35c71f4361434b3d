"""bf16 kernels (ops/act16.py, ops/mixed.py) against an INDEPENDENT oracle: torch.nn.functional in float64 with
autograd on the bf16-rounded inputs -- not the modules' own CPU branches (a padding / geometry / formula bug shared by
a module's CPU and GPU branches would pass a self-comparison)."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
BF = torch.bfloat16
DEV = "cuda"


def _frac_close(got, want, rtol, atol, what):
    got, want = got.detach().double().cpu(), want.detach().double().cpu()
    assert got.shape == want.shape, (what, got.shape, want.shape)
    bad = ((got - want).abs() > atol + rtol * want.abs()).double().mean().item()
    assert bad < 2e-3, f"{what}: {bad:.3%} of elements out of tolerance, max err {(got - want).abs().max().item():.4g}"


@pytest.mark.parametrize("M,C,relu", [(6272, 64, True), (1568, 256, False), (777, 128, True)])
def test_bn_bf16_against_functional_batch_norm(M, C, relu):
    from tfde_b200.ops import act16
    g = torch.Generator().manual_seed(M + C)
    x = (torch.randn(M, C, generator=g) * 1.5 + 0.3).to(BF)
    dy = torch.randn(M, C, generator=g).to(BF)
    gamma = torch.rand(C, generator=g) + 0.5
    beta = torch.randn(C, generator=g) * 0.1
    # oracle: F.batch_norm (training, biased variance, eps 1e-3) + ReLU in float64, gradients by autograd
    xd = x.double().requires_grad_(True)
    gd, bd = gamma.double().requires_grad_(True), beta.double().requires_grad_(True)
    rm, rv = torch.zeros(C, dtype=torch.float64), torch.ones(C, dtype=torch.float64)
    y = F.batch_norm(xd, rm, rv, gd, bd, training=True, momentum=0.01, eps=1e-3)
    if relu:
        y = F.relu(y)
    y.backward(dy.double())
    mm, mv = torch.zeros(C, device=DEV), torch.ones(C, device=DEV)
    yg, sm, si, sc, sh = act16.bn_fwd(x.to(DEV), gamma.to(DEV), beta.to(DEV), mm, mv, 0.99, 1e-3, True, relu, True)
    dxg, dgg, dbg = act16.bn_bwd(x.to(DEV), dy.to(DEV), gamma.to(DEV), sm, si, sc, sh, relu)
    _frac_close(yg, y, 2e-2, 2e-2, "bn y")                       # one bf16 ulp of the stored output
    _frac_close(mm, rm, 1e-3, 1e-5, "moving mean")
    _frac_close(mv, rv, 2e-3, 1e-5, "moving variance (unbiased, like F.batch_norm)")
    _frac_close(dbg, bd.grad, 5e-3, 0.1, "dbeta")
    _frac_close(dgg, gd.grad, 5e-3, 0.1, "dgamma")
    _frac_close(dxg, xd.grad, 3e-2, 3e-2, "bn dx")


@pytest.mark.parametrize("M,K,N,relu", [(1024, 5408, 64, True), (256, 64, 128, False), (384, 1176, 200, False)])
def test_mixed_dense_against_float64_matmul(M, K, N, relu):
    from tfde_b200.ops import mixed
    g = torch.Generator().manual_seed(K)
    x, w = torch.randn(M, K, generator=g), torch.randn(K, N, generator=g) * 0.05
    b, dy = torch.randn(N, generator=g), torch.randn(M, N, generator=g)
    xd = x.to(BF).double().requires_grad_(True)
    wd = w.to(BF).double().requires_grad_(True)
    bd = b.double().requires_grad_(True)
    y = xd @ wd + bd
    if relu:
        y = F.relu(y)
    act = "relu" if relu else "none"
    yg = mixed.dense_fwd(x.to(DEV), w.to(DEV), b.to(DEV), act)
    _frac_close(yg, y, 3e-3, 3e-3, "dense y")
    # the backward GEMMs see dy rounded to bf16 (masked by the ReLU of the module's own output)
    dym = dy.to(BF).double() * ((yg.detach().double().cpu() > 0) if relu else 1.0)
    (xd @ wd + bd).backward(dym)
    dx, dw, db = mixed.dense_bwd(x.to(DEV), w.to(DEV), yg, dy.to(DEV), act)
    _frac_close(dx, xd.grad, 5e-3, 5e-3, "dense dx")
    _frac_close(dw, wd.grad, 5e-3, 3e-2, "dense dw")
    _frac_close(db, (dy.double() * ((yg.detach().double().cpu() > 0) if relu else 1.0)).sum(0), 1e-3, 1e-2, "dense db")


@pytest.mark.parametrize("B,H,C,k,s,Co", [(8, 14, 64, 3, 1, 64), (4, 28, 64, 1, 2, 128), (2, 30, 16, 7, 2, 64)])
def test_conv_bf16_against_functional_conv2d(B, H, C, k, s, Co):
    """act16.conv2d_fwd / conv2d_bwd (explicit-operand tcgen05 GEMMs) vs F.conv2d in float64 with TF 'same' padding
    computed independently here."""
    from tfde_b200.ops import act16
    g = torch.Generator().manual_seed(B * H + C)
    x = torch.randn(B, H, H, C, generator=g).to(BF)
    w = (torch.randn(k, k, C, Co, generator=g) * 0.1)
    out = -(-H // s)
    pad_total = max((out - 1) * s + k - H, 0)
    lo, hi = pad_total // 2, pad_total - pad_total // 2
    if not act16.conv_supported(tuple(x.shape), tuple(w.shape), out, out):
        pytest.skip("shape not on the bf16 conv path")
    xd = x.double().permute(0, 3, 1, 2).requires_grad_(True)
    wd = w.to(BF).double().permute(3, 2, 0, 1).requires_grad_(True)
    y = F.conv2d(F.pad(xd, (lo, hi, lo, hi)), wd, stride=s)
    dy = torch.randn(B, out, out, Co, generator=g).to(BF)
    y.backward(dy.double().permute(0, 3, 1, 2))
    yg = act16.conv2d_fwd(x.to(DEV), w.to(DEV), None, (s, s), (lo, hi, lo, hi))
    _frac_close(yg.float(), y.permute(0, 2, 3, 1), 2e-2, 2e-2, "conv y")
    dx, dw, _ = act16.conv2d_bwd(x.to(DEV), w.to(DEV), dy.to(DEV), (s, s), (lo, hi, lo, hi), True, False)
    _frac_close(dx.float(), xd.grad.permute(0, 2, 3, 1), 3e-2, 3e-2, "conv dx")
    _frac_close(dw.float(), wd.grad.permute(2, 3, 1, 0), 2e-2, 5e-2, "conv dw")
