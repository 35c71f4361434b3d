"""bf16 tensor-core dense path (tcgen05 GEMM + glue kernels) against the CPU restatement that rounds the
same operands to bf16, and a mixed-precision training run of Model A at a tensor-core-sized batch."""
import numpy as np
import pytest
import torch

import tfde_b200 as tf
from tfde_b200.ops import mixed, native
from tfde_b200.parallel.strategy import Strategy

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _rand(*shape, seed=0):
    return torch.randn(*shape, generator=torch.Generator().manual_seed(seed), dtype=torch.float32)


def _close(a, b, rtol, what):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    err = (a - b).abs().max().item()
    scale = b.abs().max().item() + 1e-12
    assert err <= rtol * scale, f"{what}: max abs err {err:.3e} (scale {scale:.3e})"


@pytest.mark.parametrize("M,N,K,bias,relu,dt", [(128, 64, 64, False, False, torch.float32), (100, 40, 72, True, True, torch.float32),
                                                (1024, 64, 5408, True, True, torch.float32), (333, 200, 1176, False, False, torch.bfloat16),
                                                (512, 512, 256, True, False, torch.bfloat16), (2048, 2048, 512, False, True, torch.float32),
                                                (576, 64, 12544, True, True, torch.float32), (200, 40, 4096, False, False, torch.bfloat16)])
def test_tcgen05_gemm(M, N, K, bias, relu, dt):
    a = (_rand(M, K, seed=1) * 0.5).to(torch.bfloat16).to(DEV)
    b = (_rand(N, K, seed=2) * 0.5).to(torch.bfloat16).to(DEV)
    bv = _rand(N, seed=3).to(DEV) if bias else None
    c = native.gemm_bf16_tn(a, b, bv, relu, dt)
    ref = a.float() @ b.float().t()
    if bias:
        ref = ref + bv
    if relu:
        ref = torch.relu(ref)
    _close(c.float(), ref, 2e-3 if dt == torch.float32 else 1.2e-2, "tcgen05 gemm")


def test_glue_kernels():
    x, m = _rand(257, 70, seed=1), _rand(257, 70, seed=2)
    y = mixed.cast_bf16_masked(x.to(DEV), m.to(DEV))
    assert torch.equal(y.cpu(), (x * (m > 0)).to(torch.bfloat16))
    t = native.cast_transpose_bf16(x.to(DEV), m.to(DEV))
    assert torch.equal(t.cpu(), (x * (m > 0)).t().contiguous().to(torch.bfloat16))
    s = mixed.colsum_masked(x.to(DEV), m.to(DEV))
    _close(s, (x.double() * (m > 0)).sum(0), 1e-5, "colsum")


@pytest.mark.parametrize("M,K,N,relu", [(1024, 5408, 64, True), (256, 64, 128, False)])
def test_mixed_dense_fwd_bwd_matches_bf16_rounded_reference(M, K, N, relu):
    x, w, b, dy = _rand(M, K, seed=1), _rand(K, N, seed=2) * 0.05, _rand(N, seed=3), _rand(M, N, seed=4)
    act = "relu" if relu else "none"
    y_ref = mixed.dense_fwd(x, w, b, act)
    dx_ref, dw_ref, db_ref = mixed.dense_bwd(x, w, y_ref, dy, act)
    y = mixed.dense_fwd(x.to(DEV), w.to(DEV), b.to(DEV), act)
    _close(y, y_ref, 2e-3, "mixed fwd")
    dx, dw, db = mixed.dense_bwd(x.to(DEV), w.to(DEV), y, dy.to(DEV), act)
    _close(dx, dx_ref, 3e-3, "mixed dx")
    _close(dw, dw_ref, 3e-3, "mixed dw")
    _close(db, db_ref, 1e-4, "mixed db")


def test_model_a_trains_under_mixed_bfloat16():
    tf.keras.mixed_precision.set_global_policy("mixed_bfloat16")
    try:
        g = torch.Generator().manual_seed(0)
        x = torch.rand(256, 28, 28, 1, generator=g)
        y = torch.randint(0, 10, (256,), generator=g)
        res = {}
        for dev in ("cpu", DEV):
            with Strategy(device=dev).scope():
                m = tf.models.build_and_compile_cnn_model(learning_rate=0.05, seed=2)
            losses = []
            for _ in range(5):
                m.train_on_batch(x, y)
                losses.append(m.last_step_loss())
            res[dev] = losses
            if dev == DEV:
                assert m.layers[3]._mixed, "Dense(64) must have taken the tensor-core path"
        np.testing.assert_allclose(res[DEV], res["cpu"], rtol=2e-2, atol=2e-3)
        assert res[DEV][-1] < res[DEV][0]
    finally:
        tf.keras.mixed_precision.set_global_policy("float32")


CONVS = [  # B,H,W,C, k, stride, padding, Cout, bias, relu
    (8, 16, 16, 64, 1, 1, "valid", 64, True, True),
    (8, 16, 16, 64, 3, 1, "same", 64, True, False),
    (8, 16, 16, 64, 1, 2, "valid", 128, True, False),
    (8, 32, 32, 3, 7, 2, "same", 64, True, False),
    (8, 14, 14, 32, 3, 2, "same", 72, False, True),
]


@pytest.mark.parametrize("B,H,W,C,k,s,pad,Co,bias,relu", CONVS)
def test_mixed_conv_fwd_bwd(B, H, W, C, k, s, pad, Co, bias, relu):
    from tfde_b200 import ops
    from tfde_b200.ops import mixed_conv
    x, w = _rand(B, H, W, C, seed=1), _rand(k, k, C, Co, seed=2) * 0.1
    b = _rand(Co, seed=3) if bias else None
    ho, wo, pads = ops.conv_geometry(H, W, k, k, (s, s), pad)
    assert mixed_conv.supported(x.shape, w.shape, ho, wo)
    dy = _rand(B, ho, wo, Co, seed=4)
    y_ref = mixed_conv.conv2d_fwd(x, w, b, (s, s), pads, relu)
    dx_ref, dw_ref, db_ref = mixed_conv.conv2d_bwd(x, w, y_ref, dy, (s, s), pads, relu, True, bias)
    y = mixed_conv.conv2d_fwd(x.to(DEV), w.to(DEV), b.to(DEV) if bias else None, (s, s), pads, relu)
    _close(y, y_ref, 3e-3, "mixed conv fwd")
    dx, dw, db = mixed_conv.conv2d_bwd(x.to(DEV), w.to(DEV), y, dy.to(DEV), (s, s), pads, relu, True, bias)
    _close(dx, dx_ref, 4e-3, "mixed conv dx")
    _close(dw, dw_ref, 4e-3, "mixed conv dw")
    if bias:
        _close(db, db_ref, 1e-4, "mixed conv db")
