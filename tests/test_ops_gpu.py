"""Numerics of every sm_100a kernel against the plain-PyTorch fp32/fp64 restatement of the same op
(tier "Op numerics, GPU"). Includes odd batch sizes, B=1, shapes that are not a
multiple of any tile, [B] and [B,1] labels."""
import numpy as np
import pytest
import torch

from tfde_b200 import ops
from tfde_b200.ops import native, reference as ref

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _close(a, b, rtol=2e-4, atol=2e-5, what=""):
    a = a.detach().double().cpu()
    b = b.detach().double().cpu()
    err = (a - b).abs().max().item()
    scale = b.abs().max().item() + 1e-12
    assert a.shape == b.shape, f"{what}: shape {a.shape} vs {b.shape}"
    assert err <= atol + rtol * scale, f"{what}: max abs err {err:.3e} (scale {scale:.3e})"


def _rand(*shape, seed=0):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(*shape, generator=g, dtype=torch.float32)


def test_native_library_loaded():
    L = native.load()
    assert L is not None and native.available()


@pytest.mark.parametrize("M,K,N,act", [(64, 5408, 64, "relu"), (64, 64, 10, "none"), (1, 17, 5, "none"),
                                       (37, 1176, 200, "none"), (128, 200, 10, "softmax"), (257, 300, 70, "relu")])
def test_dense_fwd_bwd(M, K, N, act):
    x, w, b = _rand(M, K, seed=1), _rand(K, N, seed=2) * 0.05, _rand(N, seed=3)
    dy = _rand(M, N, seed=4)
    y_ref = ref.dense_fwd(x.double(), w.double(), b.double(), act)
    dx_ref, dw_ref, db_ref = ref.dense_bwd(x.double(), w.double(), y_ref, dy.double(), act)
    xd, wd, bd, dyd = x.to(DEV), w.to(DEV), b.to(DEV), dy.to(DEV)
    y = ops.dense_fwd(xd, wd, bd, act)
    _close(y, y_ref, what="dense fwd")
    dx, dw, db = ops.dense_bwd(xd, wd, y, dyd, act)
    _close(dx, dx_ref, what="dense dx")
    _close(dw, dw_ref, what="dense dw", rtol=5e-4)
    _close(db, db_ref, what="dense db", rtol=5e-4)


def test_dense_no_bias_no_dx():
    x, w, dy = _rand(33, 50, seed=1), _rand(50, 20, seed=2), _rand(33, 20, seed=3)
    y = ops.dense_fwd(x.to(DEV), w.to(DEV), None, "none")
    _close(y, x.double() @ w.double(), what="fwd")
    dx, dw, db = ops.dense_bwd(x.to(DEV), w.to(DEV), y, dy.to(DEV), "none", need_dx=False, has_bias=False)
    assert dx is None and db is None
    _close(dw, x.double().t() @ dy.double(), what="dw")


CONVS = [  # B,H,W,C, k, stride, padding, Cout, bias, relu
    (4, 28, 28, 1, 3, 1, "valid", 32, True, True),
    (5, 28, 28, 1, 3, 1, "same", 6, False, False),
    (3, 28, 28, 6, 6, 2, "same", 12, False, False),
    (3, 14, 14, 12, 6, 2, "same", 24, False, False),
    (1, 9, 7, 3, 3, 2, "same", 5, True, True),
    (2, 16, 16, 16, 1, 1, "valid", 64, False, False),
    (2, 15, 15, 8, 7, 2, "same", 16, True, False),
]


@pytest.mark.parametrize("B,H,W,C,k,s,pad,Co,bias,relu", CONVS)
def test_conv2d_fwd_bwd(B, H, W, C, k, s, pad, Co, bias, relu):
    x, w = _rand(B, H, W, C, seed=1), _rand(k, k, C, Co, seed=2) * 0.2
    b = _rand(Co, seed=3) if bias else None
    ho, wo, pads = ops.conv_geometry(H, W, k, k, (s, s), pad)
    dy = _rand(B, ho, wo, Co, seed=4)
    y_ref = ref.conv2d_fwd(x.double(), w.double(), b.double() if bias else None, (s, s), pads, relu)
    dx_ref, dw_ref, db_ref = ref.conv2d_bwd(x.double(), w.double(), y_ref, dy.double(), (s, s), pads, relu,
                                            True, bias)
    y = ops.conv2d_fwd(x.to(DEV), w.to(DEV), b.to(DEV) if bias else None, (s, s), pads, relu)
    _close(y, y_ref, what="conv fwd")
    dx, dw, db = ops.conv2d_bwd(x.to(DEV), w.to(DEV), y, dy.to(DEV), (s, s), pads, relu, True, bias)
    _close(dx, dx_ref, what="conv dx", rtol=5e-4)
    _close(dw, dw_ref, what="conv dw", rtol=5e-4)
    if bias:
        _close(db, db_ref, what="conv db", rtol=5e-4)


SMALLCONVS = [  # B,H,W,C, k, stride, padding, Cout, need_dx   (no bias, no activation: Model B's convolutions)
    (128, 28, 28, 1, 3, 1, "same", 6, False),
    (128, 28, 28, 6, 6, 2, "same", 12, True),
    (128, 14, 14, 12, 6, 2, "same", 24, True),
    (2, 11, 9, 4, 5, 2, "same", 8, True),
    (3, 13, 13, 2, 3, 1, "valid", 4, True),
    (1, 10, 12, 6, 4, 3, "same", 10, True),
    (5, 7, 7, 8, 3, 1, "same", 2, True),
]


@pytest.mark.parametrize("B,H,W,C,k,s,pad,Co,need_dx", SMALLCONVS)
def test_small_channel_direct_conv(B, H, W, C, k, s, pad, Co, need_dx):
    """csrc/kernels/small_conv.cu against the fp64 PyTorch restatement, and proof that it is the path that ran."""
    from tfde_b200.ops import native
    x, w = _rand(B, H, W, C, seed=11), _rand(k, k, C, Co, seed=12) * 0.2
    ho, wo, pads = ops.conv_geometry(H, W, k, k, (s, s), pad)
    dy = _rand(B, ho, wo, Co, seed=13)
    y_ref = ref.conv2d_fwd(x.double(), w.double(), None, (s, s), pads, False)
    dx_ref, dw_ref, _ = ref.conv2d_bwd(x.double(), w.double(), y_ref, dy.double(), (s, s), pads, False, need_dx, False)
    sink = []
    prev = native.set_trace_sink(sink)
    try:
        y = ops.conv2d_fwd(x.to(DEV), w.to(DEV), None, (s, s), pads, False)
        dx, dw, _ = ops.conv2d_bwd(x.to(DEV), w.to(DEV), y, dy.to(DEV), (s, s), pads, False, need_dx, False)
    finally:
        native.set_trace_sink(prev)
    assert sink == ["tfde_smallconv_fwd", "tfde_smallconv_bwd"], sink
    _close(y, y_ref, what="conv fwd")
    _close(dw, dw_ref, what="conv dw", rtol=5e-4)
    if need_dx:
        _close(dx, dx_ref, what="conv dx", rtol=5e-4)
    else:
        assert dx is None


@pytest.mark.parametrize("B,u8", [(64, False), (3, False), (1, True), (65, True)])
def test_fused_conv_relu_pool(B, u8):
    g = torch.Generator().manual_seed(5)
    if u8:
        x = torch.randint(0, 256, (B, 28, 28, 1), generator=g, dtype=torch.uint8)
        xf = x.double() / 255.0
    else:
        x = torch.rand(B, 28, 28, 1, generator=g)
        xf = x.double()
    w, b = _rand(3, 3, 1, 32, seed=6) * 0.3, _rand(32, seed=7) * 0.1
    p_ref, i_ref = ref.conv_relu_pool_fwd(xf, w.double(), b.double())
    p, i = ops.conv_relu_pool_fwd(x.to(DEV), w.to(DEV), b.to(DEV))
    _close(p, p_ref, what="pooled")
    dp = _rand(B, 13, 13, 32, seed=8)
    # use the device arg-max for the reference backward (ties between equal maxima may differ)
    _, dw_ref, db_ref = ref.conv_relu_pool_bwd(xf, w.double(), p_ref, i.cpu(), dp.double())
    _, dw, db = ops.conv_relu_pool_bwd(x.to(DEV), w.to(DEV), p, i, dp.to(DEV))
    _close(dw, dw_ref, what="fused dw", rtol=5e-4)
    _close(db, db_ref, what="fused db", rtol=5e-4)
    # arg-max agrees wherever the pooled value is positive and unique
    pos = (p_ref > 1e-6)
    assert (i.cpu()[pos] == i_ref[pos]).float().mean() > 0.999


@pytest.mark.parametrize("B,H,W,C,pool,stride,pad", [(4, 26, 26, 32, 2, 2, "valid"), (2, 13, 11, 5, 3, 2, "same"),
                                                      (1, 7, 7, 3, 2, 2, "valid")])
def test_maxpool(B, H, W, C, pool, stride, pad):
    x = _rand(B, H, W, C, seed=1)
    ho, wo, pads = ops.conv_geometry(H, W, pool, pool, (stride, stride), pad)
    y_ref, i_ref = ref.maxpool_fwd(x.double(), (pool, pool), (stride, stride), pads)
    y, i = ops.maxpool_fwd(x.to(DEV), (pool, pool), (stride, stride), pads)
    _close(y, y_ref, what="maxpool y")
    assert torch.equal(i.cpu(), i_ref)
    dy = _rand(B, ho, wo, C, seed=2)
    dx_ref = ref.maxpool_bwd(dy.double(), i_ref, x.shape, (pool, pool), (stride, stride), pads)
    dx = ops.maxpool_bwd(dy.to(DEV), i, tuple(x.shape), (pool, pool), (stride, stride), pads)
    _close(dx, dx_ref, what="maxpool dx")


@pytest.mark.parametrize("shape,gamma,relu", [((128, 28, 28, 6), False, True), ((16, 7, 7, 24), False, True),
                                              ((128, 200), False, True), ((33, 5, 5, 70), True, False),
                                              ((2, 3), True, False), ((128, 14, 14, 12), False, True),
                                              ((37, 9, 11, 7), True, True), ((64, 13, 13, 32), True, False),
                                              ((128, 7, 7, 24), False, True)])
def test_batchnorm(shape, gamma, relu):
    c = shape[-1]
    x = _rand(*shape, seed=1) * 2.0 + 0.5
    gm = (_rand(c, seed=2) * 0.1 + 1.0) if gamma else None
    bt = _rand(c, seed=3) * 0.1
    mm, mv = _rand(c, seed=4) * 0.1, torch.rand(c) + 0.5
    dy = _rand(*shape, seed=5)
    unb = len(shape) == 4
    y_ref, sm_ref, si_ref, mm_ref, mv_ref = ref.bn_fwd(x.double(), gm.double() if gamma else None, bt.double(),
                                                       mm.double(), mv.double(), 0.99, 1e-3, True, relu, unb)
    dx_ref, dg_ref, db_ref = ref.bn_bwd(x.double(), y_ref, dy.double(), gm.double() if gamma else None, sm_ref,
                                        si_ref, relu)
    mmd, mvd = mm.to(DEV), mv.to(DEV)
    y, sm, si = ops.bn_fwd(x.to(DEV), gm.to(DEV) if gamma else None, bt.to(DEV), mmd, mvd, 0.99, 1e-3, True, relu, unb)
    _close(y, y_ref, what="bn y", rtol=1e-3, atol=1e-4)
    _close(sm, sm_ref, what="bn mean", rtol=1e-4)
    _close(si, si_ref, what="bn invstd", rtol=1e-3)
    _close(mmd, mm_ref, what="moving mean")
    _close(mvd, mv_ref, what="moving var", rtol=1e-3)
    dx, dg, db = ops.bn_bwd(x.to(DEV), y, dy.to(DEV), gm.to(DEV) if gamma else None, sm, si, relu)
    _close(dx, dx_ref, what="bn dx", rtol=2e-3, atol=1e-4)
    _close(db, db_ref, what="bn dbeta", rtol=1e-3)
    if gamma:
        _close(dg, dg_ref, what="bn dgamma", rtol=1e-3)
    # inference path with moving statistics
    yi_ref = ref.bn_fwd(x.double(), gm.double() if gamma else None, bt.double(), mm_ref, mv_ref, 0.99, 1e-3,
                        False, relu, unb)[0]
    yi, _, _ = ops.bn_fwd(x.to(DEV), gm.to(DEV) if gamma else None, bt.to(DEV), mmd, mvd, 0.99, 1e-3, False, relu, unb)
    _close(yi, yi_ref, what="bn infer", rtol=1e-3, atol=1e-4)


def test_elementwise_and_softmax():
    x = _rand(77, 13, seed=1)
    _close(ops.relu_fwd(x.to(DEV)), torch.relu(x), what="relu")
    y = torch.relu(x)
    dy = _rand(77, 13, seed=2)
    _close(ops.relu_bwd(y.to(DEV), dy.to(DEV)), ref.relu_bwd(y, dy), what="relu bwd")
    p_ref = ref.softmax_fwd(x.double())
    p = ops.softmax_fwd(x.to(DEV))
    _close(p, p_ref, what="softmax")
    _close(ops.softmax_bwd(p, dy.to(DEV)), ref.softmax_bwd(p_ref, dy.double()), what="softmax bwd")
    a, b = _rand(1001, seed=3), _rand(1001, seed=4)
    _close(ops.add(a.to(DEV), b.to(DEV), relu=True), torch.relu(a + b), what="add relu")
    u = torch.randint(0, 256, (513,), dtype=torch.uint8)
    _close(ops.u8_to_f32(u.to(DEV)), u.float() / 255.0, what="u8")
    g = _rand(3, 5, 4, 9, seed=5)
    _close(ops.global_avgpool_fwd(g.to(DEV)), ref.global_avgpool_fwd(g), what="gap")
    dg = _rand(3, 9, seed=6)
    _close(ops.global_avgpool_bwd(dg.to(DEV), tuple(g.shape)), ref.global_avgpool_bwd(dg, g.shape), what="gap bwd")


@pytest.mark.parametrize("n,rate", [(128 * 200, 0.5), (1003, 0.25), (5, 0.9)])
def test_dropout_matches_philox_reference(n, rate):
    x = _rand(n, seed=1)
    seed, step, stream = 0x1234567890AB, 7, 3
    hp = native.make_hparams(DEV, seed=seed, step=step)
    y = ops.dropout(x.to(DEV), rate, seed, step, stream, hparams=hp, training=True)
    y_ref = ref.dropout_fwd(x, rate, seed, step, stream)
    assert torch.equal(y.cpu() != 0, y_ref != 0), "dropout mask differs from the Philox reference"
    _close(y, y_ref, what="dropout")
    assert torch.equal(ops.dropout(x.to(DEV), rate, seed, step, stream, hparams=hp, training=False).cpu(), x)


@pytest.mark.parametrize("rows,C,probs,lshape", [(64, 10, False, "flat"), (128, 10, True, "col"), (1, 3, False, "flat"),
                                                 (1000, 37, False, "col"), (333, 10, True, "flat")])
def test_xent(rows, C, probs, lshape):
    z = _rand(rows, C, seed=1) * 3
    g = torch.Generator().manual_seed(2)
    lab = torch.randint(0, C, (rows,), generator=g)
    labels = lab if lshape == "flat" else lab.reshape(-1, 1)
    scale = 1.0 / (rows * 2)
    inp = torch.softmax(z, -1) if probs else z
    fn = ref.xent_probs if probs else ref.xent_logits
    l_ref, c_ref, d_ref = fn(inp.double(), labels, scale)
    o4, d = ops.xent(inp.to(DEV), labels.to(DEV), scale, probs)
    o4 = o4.cpu()
    assert abs(o4[0].item() - l_ref.item()) <= 1e-4 * max(1.0, abs(l_ref.item()))
    assert o4[1].item() == c_ref.item() and o4[2].item() == rows
    _close(d, d_ref, what="dloss")
    # int32 labels too
    o4b, _ = ops.xent(inp.to(DEV), lab.to(torch.int32).to(DEV), scale, probs)
    assert torch.allclose(o4b.cpu(), o4)


@pytest.mark.parametrize("probs", [False, True])
def test_xent_out_of_range_label_is_nan_loss_zero_gradient(probs):
    """No out-of-bounds read for a label outside [0, C): NaN loss, zero gradient for that row, other rows intact."""
    rows, C = 37, 10
    z = _rand(rows, C, seed=3) * 2
    lab = torch.randint(0, C, (rows,), generator=torch.Generator().manual_seed(4))
    lab[5], lab[20] = C + 2, -1
    inp = torch.softmax(z, -1) if probs else z
    fn = ref.xent_probs if probs else ref.xent_logits
    l_ref, c_ref, d_ref = fn(inp.double(), lab, 0.5)
    o4, d = ops.xent(inp.to(DEV), lab.to(DEV), 0.5, probs)
    assert torch.isnan(o4[0].cpu()) and torch.isnan(l_ref)
    assert o4[1].item() == c_ref.item()
    assert float(d[5].abs().max()) == 0.0 and float(d[20].abs().max()) == 0.0
    _close(d, d_ref, what="dloss")


@pytest.mark.parametrize("opt", ["sgd", "momentum", "nesterov", "adam"])
def test_optimizer_apply(opt):
    from tfde_b200.parallel.comm import Arena
    from tfde_b200.keras import optimizers as O
    n = 10007
    mk = {"sgd": lambda: O.SGD(0.1), "momentum": lambda: O.SGD(0.1, momentum=0.9),
          "nesterov": lambda: O.SGD(0.1, momentum=0.9, nesterov=True), "adam": lambda: O.Adam(0.01)}[opt]
    w0, grads = _rand(n, seed=1), [_rand(n, seed=10 + i) for i in range(3)]
    res = {}
    for dev in ("cpu", DEV):
        o = mk().bind(dev, seed=1)
        a = Arena(n, dev)
        a.w[:n] = w0.to(dev)
        for i, g in enumerate(grads):
            a.g[:n] = g.to(dev)
            a.g[n:n + 4] = torch.tensor([0.5 + i, 3.0, 64.0, 0.0], device=dev)
            o.apply_local(a)
        res[dev] = (a.w.cpu().clone(), a.metrics.cpu().clone())
        if dev == DEV:
            assert native.hparams_get(o.hparams)["step"] == 3
    _close(res[DEV][0][:n], res["cpu"][0][:n], rtol=1e-4, atol=1e-5, what=f"{opt} weights")
    _close(res[DEV][1], res["cpu"][1], what="metrics")
    assert res[DEV][1][0].item() == pytest.approx(0.5 + 1.5 + 2.5)
