"""Latency kernels for the reference's batch-64 operating point: cluster/DSMEM split-K dense and the
fused classifier head, against the PyTorch restatement."""
import numpy as np
import pytest
import torch

from tfde_b200 import ops
from tfde_b200.ops import native, reference as ref

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _rand(*shape, seed=0):
    return torch.randn(*shape, generator=torch.Generator().manual_seed(seed), dtype=torch.float32)


def _close(a, b, rtol=2e-4, atol=2e-5, what=""):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    err = (a - b).abs().max().item()
    scale = b.abs().max().item() + 1e-12
    assert err <= atol + rtol * scale, f"{what}: max abs err {err:.3e} (scale {scale:.3e})"


@pytest.mark.parametrize("M,K,N,relu,bias", [(64, 5408, 64, True, True), (1, 5408, 64, False, True),
                                             (37, 1176, 200, True, False), (128, 1176, 200, False, False), (128, 2048, 128, False, True),
                                             (64, 512, 4, True, True), (50, 16000, 60, False, False)])
def test_dense_skinny_cluster_splitk(M, K, N, relu, bias):
    L = native.load()
    assert L.tfde_dense_skinny_supported(M, K, N) == 1
    x, w = _rand(M, K, seed=1), _rand(K, N, seed=2) * 0.05
    b = _rand(N, seed=3) if bias else None
    y_ref = ref.dense_fwd(x.double(), w.double(), b.double() if bias else None, "relu" if relu else "none")
    xd, wd, bd = x.to(DEV), w.to(DEV), (b.to(DEV) if bias else None)
    for _ in range(2):   # twice: tickets must be self-cleaning
        y = ops.dense_fwd(xd, wd, bd, "relu" if relu else "none")
        _close(y, y_ref, what="skinny dense")
    # bit-identical to itself across launches (fixed reduction order)
    assert torch.equal(ops.dense_fwd(xd, wd, bd, "relu" if relu else "none"), y)


@pytest.mark.parametrize("M,K,N,probs,bias,lab", [(64, 64, 10, False, True, "i64"), (128, 200, 10, True, True, "i64"),
                                                  (1, 64, 10, False, True, "i32"), (333, 64, 10, False, False, "i64"),
                                                  (1000, 37, 32, True, True, "i64"), (64, 5, 2, False, True, "i64")])
def test_fused_head(M, K, N, probs, bias, lab):
    h = torch.relu(_rand(M, K, seed=1))
    w, b = _rand(K, N, seed=2) * 0.3, (_rand(N, seed=3) if bias else None)
    labels = torch.randint(0, N, (M,), generator=torch.Generator().manual_seed(4))
    scale = 1.0 / (2 * M)
    act = "softmax" if probs else "none"
    z_ref = ref.dense_fwd(h.double(), w.double(), b.double() if bias else None, act)
    fn = ref.xent_probs if probs else ref.xent_logits
    l_ref, c_ref, d_ref = fn(z_ref, labels, scale)
    dx_ref, dw_ref, db_ref = ref.dense_bwd(h.double(), w.double(), z_ref, d_ref, act, True, bias)
    ld = labels.to(torch.int32 if lab == "i32" else torch.int64).to(DEV)
    out = torch.empty(M, N, device=DEV)
    o4, out, dh, dw, db = ops.head_fwd_bwd(h.to(DEV), w.to(DEV), b.to(DEV) if bias else None, ld, scale, probs,
                                           False, True, out=out)
    o4 = o4.cpu()
    assert abs(o4[0].item() - l_ref.item()) <= 2e-4 * max(1.0, abs(l_ref.item()))
    assert o4[1].item() == c_ref.item() and o4[2].item() == M
    _close(out, z_ref, what="head out", rtol=1e-4)
    _close(dh, dx_ref, what="head dh", rtol=5e-4)
    _close(dw, dw_ref, what="head dw", rtol=5e-4)
    if bias:
        _close(db, db_ref, what="head db", rtol=5e-4)
    # relu_mask variant
    _, _, dh2, _, _ = ops.head_fwd_bwd(h.to(DEV), w.to(DEV), b.to(DEV) if bias else None, ld, scale, probs, True, True)
    _close(dh2, dx_ref * (h.double() > 0), what="head dh masked", rtol=5e-4)


@pytest.mark.parametrize("M,K,relu,need_dx,bias", [(64, 5408, True, True, True), (33, 5408, False, True, True),
                                                   (64, 100, True, False, False), (1, 64, True, True, True)])
def test_dense_bwd_skinny_one_launch(M, K, relu, need_dx, bias):
    """dx, dW, db of Dense(K -> 64) at batch <= 64 in one launch vs the fp32 PyTorch restatement (partial last
    feature tile: 5408 = 84 * 64 + 32; partial batch; no data gradient; no bias)."""
    L = native.load()
    N = 64
    assert L.tfde_dense_bwd_skinny_supported(M, K, N) == 1
    assert L.tfde_dense_bwd_skinny_supported(65, K, N) == 0 and L.tfde_dense_bwd_skinny_supported(M, K, 32) == 0
    x, w, dy = _rand(M, K, seed=1), _rand(K, N, seed=2) * 0.05, _rand(M, N, seed=3)
    y = torch.relu(x @ w) if relu else x @ w
    act = "relu" if relu else "none"
    rdx, rdw, rdb = ref.dense_bwd(x, w, y, dy, act, need_dx, bias)
    dx, dw, db = ops.dense_bwd(x.to(DEV), w.to(DEV), y.to(DEV), dy.to(DEV), act, need_dx, bias)
    torch.cuda.synchronize()
    _close(dw, rdw, what="dw")
    if bias:
        _close(db, rdb, what="db")
    if need_dx:
        _close(dx, rdx, what="dx")
    else:
        assert dx is None
