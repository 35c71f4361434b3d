"""bf16-activation kernels (csrc/kernels/act_bf16.cu, ops/act16.py) against their fp32 PyTorch restatement:
the CPU branch of every op is the plain fp32 reference applied to the same bf16-rounded inputs, rounded to
bf16 where the kernel stores bf16. Tolerances are one bf16 ulp of the result (2^-8 relative) plus the
reduction-order noise of the fp32 statistics."""
import pytest
import torch

pytestmark = pytest.mark.gpu

BF = torch.bfloat16


def _dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    return torch.device("cuda", 0)


def _close(a, b, rtol=2e-2, atol=2e-2):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    assert a.shape == b.shape
    err = (a - b).abs()
    bound = atol + rtol * b.abs()
    frac_bad = (err > bound).float().mean().item()
    assert frac_bad < 1e-3, f"{frac_bad:.4%} of elements out of tolerance, max err {err.max().item():.4g}"


@pytest.mark.parametrize("M,C,relu", [(6272, 64, True), (1568, 256, False), (392, 512, True), (200, 2048, True),
                                      (777, 128, True)])
def test_bn_fwd_bwd(M, C, relu):
    from tfde_b200.ops import act16
    dev = _dev()
    g = torch.Generator().manual_seed(M + C)
    x = (torch.randn(M, C, generator=g) * 1.5 + 0.3).to(BF)
    dy = torch.randn(M, C, generator=g).to(BF)
    gamma = torch.rand(C, generator=g) + 0.5
    beta = torch.randn(C, generator=g) * 0.1
    mm, mv = torch.zeros(C), torch.ones(C)
    yc, smc, sic, scc, shc = act16.bn_fwd(x, gamma, beta, mm, mv, 0.99, 1e-3, True, relu, True)
    dxc, dgc, dbc = act16.bn_bwd(x, dy, gamma, smc, sic, scc, shc, relu)
    mmg, mvg = torch.zeros(C, device=dev), torch.ones(C, device=dev)
    yg, smg, sig, scg, shg = act16.bn_fwd(x.to(dev), gamma.to(dev), beta.to(dev), mmg, mvg, 0.99, 1e-3, True, relu, True)
    dxg, dgg, dbg = act16.bn_bwd(x.to(dev), dy.to(dev), gamma.to(dev), smg, sig, scg, shg, relu)
    torch.cuda.synchronize()
    assert yg.dtype == BF and dxg.dtype == BF
    _close(smg, smc, 1e-4, 1e-4)
    _close(sig, sic, 1e-3, 1e-4)
    _close(mmg, mm, 1e-4, 1e-5)
    _close(mvg, mv, 1e-3, 1e-5)
    _close(yg, yc)
    _close(dbg, dbc, 2e-3, 0.05)
    _close(dgg, dgc, 2e-3, 0.05)
    _close(dxg, dxc, 3e-2, 3e-2)


def test_bn_inference_and_colsum():
    from tfde_b200.ops import act16
    dev = _dev()
    g = torch.Generator().manual_seed(1)
    x = torch.randn(1000, 64, generator=g).to(BF)
    gamma, beta = torch.rand(64, generator=g) + 0.5, torch.randn(64, generator=g)
    mm, mv = torch.randn(64, generator=g) * 0.1, torch.rand(64, generator=g) + 0.5
    yc = act16.bn_fwd(x, gamma, beta, mm.clone(), mv.clone(), 0.99, 1e-3, False, True, True)[0]
    yg = act16.bn_fwd(x.to(dev), gamma.to(dev), beta.to(dev), mm.to(dev), mv.to(dev), 0.99, 1e-3, False, True, True)[0]
    _close(yg, yc)
    _close(act16.colsum(x.to(dev)), x.float().sum(0), 1e-3, 1e-2)


def test_elementwise_and_pools():
    from tfde_b200.ops import act16
    dev = _dev()
    g = torch.Generator().manual_seed(2)
    a = torch.randn(8, 14, 14, 64, generator=g).to(BF)
    b = torch.randn(8, 14, 14, 64, generator=g).to(BF)
    _close(act16.add(a.to(dev), b.to(dev), relu=True), act16.add(a, b, relu=True), 1e-2, 1e-2)
    _close(act16.relu_bwd(a.to(dev), b.to(dev)), act16.relu_bwd(a, b), 0, 0)
    _close(act16.relu_bwd(a.to(dev), b.to(dev), addend=a.to(dev)), act16.relu_bwd(a, b, addend=a), 1e-2, 1e-2)
    x32 = torch.randn(1000, generator=g)
    _close(act16.cast_to_bf16(x32.to(dev)), x32.to(BF), 0, 0)
    # max-pool 3x3 / stride 2 / same (the ResNet stem) and its backward
    x = torch.randn(4, 17, 17, 64, generator=g).to(BF)
    pads = (0, 1, 0, 1)
    yc, ic = act16.maxpool_fwd(x, (3, 3), (2, 2), pads)
    yg, ig = act16.maxpool_fwd(x.to(dev), (3, 3), (2, 2), pads)
    _close(yg, yc, 0, 0)
    assert torch.equal(ig.cpu(), ic)
    dy = torch.randn(*yc.shape, generator=g).to(BF)
    _close(act16.maxpool_bwd(dy.to(dev), ig, tuple(x.shape), (3, 3), (2, 2), pads),
           act16.maxpool_bwd(dy, ic, tuple(x.shape), (3, 3), (2, 2), pads), 1e-2, 1e-2)
    # overlapping windows with more than 2 x 2 covers per input position (3x3 / stride 1): the general backward loop;
    # 2x2 / stride 2: exactly one cover (the batched-load path, like 3x3 / 2 above)
    for pool, st, pd in (((3, 3), (1, 1), (1, 1, 1, 1)), ((2, 2), (2, 2), (0, 0, 0, 0))):
        xs = torch.randn(2, 12, 12, 32, generator=g).to(BF)
        yc2, ic2 = act16.maxpool_fwd(xs, pool, st, pd)
        yg2, ig2 = act16.maxpool_fwd(xs.to(dev), pool, st, pd)
        _close(yg2, yc2, 0, 0)
        assert torch.equal(ig2.cpu(), ic2)
        dy2 = torch.randn(*yc2.shape, generator=g).to(BF)
        _close(act16.maxpool_bwd(dy2.to(dev), ig2, tuple(xs.shape), pool, st, pd),
               act16.maxpool_bwd(dy2, ic2, tuple(xs.shape), pool, st, pd), 1e-2, 1e-2)
    # global average pool
    _close(act16.global_avgpool_fwd(a.to(dev)), act16.global_avgpool_fwd(a), 1e-4, 1e-4)
    d = torch.randn(8, 64, generator=g)
    _close(act16.global_avgpool_bwd(d.to(dev), tuple(a.shape)), act16.global_avgpool_bwd(d, tuple(a.shape)), 1e-2, 1e-4)


@pytest.mark.parametrize("cfg", [
    dict(B=8, H=16, W=16, C=64, K=1, Co=256, s=1, pad="same"),     # 1x1: activation is the GEMM operand
    dict(B=8, H=16, W=16, C=64, K=3, Co=64, s=1, pad="same"),      # 3x3: bf16 im2col / dgrad gather
    dict(B=8, H=16, W=16, C=128, K=1, Co=256, s=2, pad="same"),    # strided 1x1 (projection shortcut)
    dict(B=8, H=32, W=32, C=3, K=7, Co=64, s=2, pad="same", f32in=True),   # stem: fp32 image in
])
def test_conv(cfg):
    from tfde_b200 import ops
    from tfde_b200.ops import act16
    dev = _dev()
    g = torch.Generator().manual_seed(cfg["K"] * 100 + cfg["C"])
    B, H, W, C, K, Co, s = (cfg[k] for k in ("B", "H", "W", "C", "K", "Co", "s"))
    ho, wo, pads = ops.conv_geometry(H, W, K, K, (s, s), cfg["pad"])
    x = torch.randn(B, H, W, C, generator=g)
    if not cfg.get("f32in"):
        x = x.to(BF)
    w = torch.randn(K, K, C, Co, generator=g) * (1.0 / (K * K * C) ** 0.5)
    bias = torch.randn(Co, generator=g) * 0.1
    dy = torch.randn(B, ho, wo, Co, generator=g).to(BF)
    need_dx = not cfg.get("f32in")
    yc = act16.conv2d_fwd(x, w, bias, (s, s), pads)
    dxc, dwc, dbc = act16.conv2d_bwd(x, w, dy, (s, s), pads, need_dx=need_dx)
    yg = act16.conv2d_fwd(x.to(dev), w.to(dev), bias.to(dev), (s, s), pads)
    dxg, dwg, dbg = act16.conv2d_bwd(x.to(dev), w.to(dev), dy.to(dev), (s, s), pads, need_dx=need_dx)
    torch.cuda.synchronize()
    assert yg.dtype == BF
    _close(yg, yc)
    _close(dwg, dwc, 1e-2, 5e-2)
    _close(dbg, dbc, 1e-3, 5e-2)
    if need_dx:
        assert dxg.dtype == BF
        _close(dxg, dxc)
    # the identically-zero bias gradient of conv -> batch-norm
    _, _, dbz = act16.conv2d_bwd(x.to(dev), w.to(dev), dy.to(dev), (s, s), pads, need_dx=False, bias_grad_zero=True)
    assert float(dbz.abs().max()) == 0.0


def test_resnet_bf16_activations_match_cpu_restatement():
    """A small ResNet trained for a few steps with bf16 activations: the CUDA path follows the CPU restatement
    of the same policy, and both follow the fp32 model."""
    import tfde_b200 as tf
    from tfde_b200.models import resnet
    from tfde_b200.parallel.strategy import Strategy
    _dev()
    g = torch.Generator().manual_seed(0)
    x = torch.rand(32, 32, 32, 3, generator=g)
    y = torch.randint(0, 10, (32,), generator=g)

    def train(device, policy, acts):
        tf.keras.mixed_precision.set_global_policy(policy, activations=acts)
        try:
            with Strategy(device=device).scope():
                m = resnet.build_resnet((1, 1, 1, 1), 16, 10, (32, 32, 3), seed=3)
                m.compile(optimizer=tf.keras.optimizers.SGD(0.05),
                          loss=tf.keras.losses.SparseCategoricalCrossentropy(from_logits=True), metrics=["accuracy"])
            out = []
            for _ in range(4):
                m.train_on_batch(x, y)
                out.append(m.last_step_loss())
            dts = [getattr(l, "_out_dtype", None) for l in m.exec_plan]
            return out, dts, m
        finally:
            tf.keras.mixed_precision.set_global_policy("float32")

    lg, dts, mg = train("cuda", "mixed_bfloat16", "bfloat16")
    lc, _, _ = train("cpu", "mixed_bfloat16", "bfloat16")
    lf, _, _ = train("cpu", "float32", None)
    assert dts[0] == BF and dts[4] == BF, dts          # stem conv and first block really ran in bf16
    assert mg.graph_captured
    assert abs(lg[0] - lc[0]) < 0.05 and abs(lg[0] - lf[0]) < 0.1, (lg, lc, lf)
    for a, b in zip(lg, lc):
        assert abs(a - b) < 0.25 * max(1.0, abs(b)), (lg, lc, lf)
    assert lg[-1] < lg[0]


@pytest.mark.parametrize("K,M,N", [(4096, 64, 64), (1000, 576, 64), (6272, 256, 128), (777, 128, 512), (50176, 64, 256)])
def test_gemm_nt_mn_major(K, M, N):
    """tcgen05 GEMM with MN-major operands (A[K,M]^T B[K,N]) against fp32 matmul of the same bf16 inputs."""
    from tfde_b200.ops import tensorcore
    dev = _dev()
    g = torch.Generator().manual_seed(K + M + N)
    a = torch.randn(K, M, generator=g).to(BF).to(dev)
    b = torch.randn(K, N, generator=g).to(BF).to(dev)
    c = tensorcore.gemm_bf16_nt(a, b)
    ref = a.float().t() @ b.float()
    torch.cuda.synchronize()
    err = (c - ref).abs().max().item()
    assert err <= 2e-3 * (K ** 0.5) * 4, f"max err {err}"
    # strided operand: columns [0, M) of a wider matrix (im2col with padded leading dimension)
    wide = torch.randn(K, M + 8, generator=g).to(BF).to(dev)
    c2 = tensorcore.gemm_bf16_nt(wide[:, :M], b)
    ref2 = wide[:, :M].float().t() @ b.float()
    assert (c2 - ref2).abs().max().item() <= 2e-3 * (K ** 0.5) * 4


@pytest.mark.parametrize("M,K,Kb,N", [(1024, 64, 64, 256), (6272, 576, 576, 64), (3000, 152, 147, 64), (512, 2048, 2048, 512)])
def test_gemm_nn_weights_as_stored(M, K, Kb, N):
    """A[M,K] K-major x B[Kb,N] row-major (MN-major): the forward GEMM on weights exactly as Keras stores them;
    Kb < K = zero-padded im2col columns meeting TMA-zero-filled weight rows."""
    from tfde_b200.ops import tensorcore
    dev = _dev()
    g = torch.Generator().manual_seed(M + K + N)
    a = torch.randn(M, K, generator=g).to(BF)
    a[:, Kb:] = 0
    a = a.to(dev)
    b = torch.randn(Kb, N, generator=g).to(BF).to(dev)
    bias = torch.randn(N, generator=g).to(dev)
    c = tensorcore.gemm_bf16_nn(a, b, bias, True, BF)
    ref = torch.relu(a[:, :Kb].float() @ b.float() + bias)
    torch.cuda.synchronize()
    _close(c, ref, 2e-2, 2e-2 * (K ** 0.5) / 8)


@pytest.mark.parametrize("B,H,C,Co", [(16, 56, 64, 256), (13, 40, 64, 64)])
def test_bn_statistics_fused_into_conv_gemm(B, H, C, Co):
    """conv (1x1) -> training batch-norm with the statistics taken from the GEMM epilogue's partial sums equals the
    two-pass batch-norm on the same convolution output (same rounded tensor, fp32/fp64 reductions both ways)."""
    from tfde_b200.ops import act16
    dev = _dev()
    g = torch.Generator().manual_seed(B + Co)
    x = torch.randn(B, H, H, C, generator=g).to(BF).to(dev)
    w = (torch.randn(1, 1, C, Co, generator=g) * 0.1).to(dev)
    bias = (torch.randn(Co, generator=g) * 0.1).to(dev)
    gamma, beta = (torch.rand(Co, generator=g) + 0.5).to(dev), (torch.randn(Co, generator=g) * 0.1).to(dev)
    saved = act16._FUSE_BN_STATS
    act16._FUSE_BN_STATS = True          # off by default (measured neutral on ResNet-50), still supported
    try:
        pshape = act16.bn_partials_shape(tuple(x.shape), tuple(w.shape), H, H)
    finally:
        act16._FUSE_BN_STATS = saved
    assert pshape is not None and pshape == ((B * H * H + 31) // 32, Co)
    psum, psq = torch.zeros(pshape, device=dev), torch.zeros(pshape, device=dev)
    y1 = act16.conv2d_fwd(x, w, bias, (1, 1), (0, 0, 0, 0), bn_partials=(psum, psq))
    y2 = act16.conv2d_fwd(x, w, bias, (1, 1), (0, 0, 0, 0))
    assert torch.equal(y1, y2)
    mm1, mv1 = torch.zeros(Co, device=dev), torch.ones(Co, device=dev)
    mm2, mv2 = torch.zeros(Co, device=dev), torch.ones(Co, device=dev)
    o1 = act16.bn_fwd(y1, gamma, beta, mm1, mv1, 0.99, 1e-3, True, True, True, partials=(psum, psq))
    o2 = act16.bn_fwd(y2, gamma, beta, mm2, mv2, 0.99, 1e-3, True, True, True)
    torch.cuda.synchronize()
    ref_sum = y1.float().reshape(-1, Co).sum(0)
    _close(psum.sum(0), ref_sum, 1e-4, 1e-2)
    for a, b in zip(o1[1:], o2[1:]):          # save_mean, save_invstd, scale, shift
        _close(a, b, 1e-5, 1e-6)
    _close(mm1, mm2, 1e-5, 1e-7)
    _close(mv1, mv2, 1e-5, 1e-7)
    _close(o1[0], o2[0], 1e-2, 1e-2)
