"""Shape fuzzing of the fp32 kernels (hypothesis) against the float64 PyTorch restatement:
dense (all three paths: generic split-K, skinny cluster kernel, one-launch backward), convolution (implicit GEMM and
the small-channel direct kernels, TF 'same'/'valid' padding, strides), batch-norm (wide and small-C reductions)."""
import pytest
import torch
from hypothesis import HealthCheck, given, settings, strategies as st

from tfde_b200 import ops
from tfde_b200.ops import reference as ref

pytestmark = pytest.mark.gpu
DEV = "cuda"
CFG = dict(max_examples=25, deadline=None, derandomize=True,
           suppress_health_check=[HealthCheck.too_slow, HealthCheck.function_scoped_fixture])


def _rand(*shape, seed):
    return torch.randn(*shape, generator=torch.Generator().manual_seed(seed))


def _close(got, want, rtol=2e-4, what=""):
    got, want = got.detach().double().cpu(), want.detach().double().cpu()
    assert got.shape == want.shape, (what, got.shape, want.shape)
    scale = max(float(want.abs().max()), 1e-6)
    err = float((got - want).abs().max()) / scale
    assert err < rtol, (what, err)


@settings(**CFG)
@given(M=st.integers(1, 130), K=st.integers(1, 700), N=st.integers(1, 80), act=st.sampled_from(["none", "relu"]),
       bias=st.booleans())
def test_fuzz_dense(M, K, N, act, bias):
    x, w, dy = _rand(M, K, seed=M), _rand(K, N, seed=K) * 0.1, _rand(M, N, seed=N)
    b = _rand(N, seed=7) if bias else None
    y_ref = ref.dense_fwd(x.double(), w.double(), b.double() if bias else None, act)
    dx_ref, dw_ref, db_ref = ref.dense_bwd(x.double(), w.double(), y_ref, dy.double(), act, True, bias)
    y = ops.dense_fwd(x.to(DEV), w.to(DEV), b.to(DEV) if bias else None, act)
    _close(y, y_ref, what="dense y")
    dx, dw, db = ops.dense_bwd(x.to(DEV), w.to(DEV), y, dy.to(DEV), act, True, bias)
    _close(dx, dx_ref, what="dense dx")
    _close(dw, dw_ref, what="dense dw")
    if bias:
        _close(db, db_ref, what="dense db")


@settings(**CFG)
@given(B=st.integers(1, 5), H=st.integers(3, 17), W=st.integers(3, 17), C=st.sampled_from([1, 2, 3, 4, 6, 8, 12, 16]),
       k=st.integers(1, 6), s=st.integers(1, 3), pad=st.sampled_from(["same", "valid"]),
       Co=st.sampled_from([2, 4, 5, 6, 8, 12, 24, 32, 40]))
def test_fuzz_conv2d(B, H, W, C, k, s, pad, Co):
    if pad == "valid" and (k > H or k > W):
        return
    x, w = _rand(B, H, W, C, seed=H * W), _rand(k, k, C, Co, seed=k + C) * 0.2
    ho, wo, pads = ops.conv_geometry(H, W, k, k, (s, s), pad)
    dy = _rand(B, ho, wo, Co, seed=Co)
    y_ref = ref.conv2d_fwd(x.double(), w.double(), None, (s, s), pads, False)
    dx_ref, dw_ref, _ = ref.conv2d_bwd(x.double(), w.double(), y_ref, dy.double(), (s, s), pads, False, True, False)
    y = ops.conv2d_fwd(x.to(DEV), w.to(DEV), None, (s, s), pads, False)
    _close(y, y_ref, what="conv y")
    dx, dw, _ = ops.conv2d_bwd(x.to(DEV), w.to(DEV), y, dy.to(DEV), (s, s), pads, False, True, False)
    _close(dx, dx_ref, rtol=5e-4, what="conv dx")
    _close(dw, dw_ref, rtol=5e-4, what="conv dw")


@settings(**CFG)
@given(M=st.integers(2, 3000), C=st.integers(1, 70), relu=st.booleans(), gamma=st.booleans())
def test_fuzz_batchnorm(M, C, relu, gamma):
    x, dy = _rand(M, C, seed=M + C) * 1.3 + 0.2, _rand(M, C, seed=C)
    gm = (torch.rand(C, generator=torch.Generator().manual_seed(3)) + 0.5) if gamma else None
    bt = _rand(C, seed=5) * 0.1
    mm, mv = torch.zeros(C), torch.ones(C)
    y_ref, sm_ref, si_ref, _, _ = ref.bn_fwd(x.double(), gm.double() if gamma else None, bt.double(),
                                             mm.double().clone(), mv.double().clone(), 0.99, 1e-3, True, relu, False)
    dx_ref, dg_ref, db_ref = ref.bn_bwd(x.double(), y_ref, dy.double(), gm.double() if gamma else None, sm_ref, si_ref, relu)
    y, sm, si = ops.bn_fwd(x.to(DEV), gm.to(DEV) if gamma else None, bt.to(DEV), mm.to(DEV), mv.to(DEV), 0.99, 1e-3,
                           True, relu, False)
    dx, dg, db = ops.bn_bwd(x.to(DEV), y, dy.to(DEV), gm.to(DEV) if gamma else None, sm, si, relu)
    _close(y, y_ref, what="bn y")
    _close(sm, sm_ref, what="bn mean")
    _close(dx, dx_ref, rtol=1e-3, what="bn dx")
    _close(db, db_ref, rtol=1e-3, what="bn dbeta")
    if gamma:
        _close(dg, dg_ref, rtol=1e-3, what="bn dgamma")
