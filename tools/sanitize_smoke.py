"""Small program for compute-sanitizer (memcheck / racecheck): three eager training steps of Model A and
Model B (every fp32 kernel incl. the cluster split-K dense and the fused head), plus -- with --tc -- a few
tcgen05 GEMMs and a mixed-precision convolution."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["TFDE_CUDA_GRAPH"] = "0"
import torch

import tfde_b200 as tf

g = torch.Generator().manual_seed(0)
x = torch.rand(32, 28, 28, 1, generator=g)
y = torch.randint(0, 10, (32,), generator=g)
ma = tf.models.build_and_compile_cnn_model(0.05, seed=1)
mb = tf.models.build_and_compile_model_b(0.05, seed=1)
for _ in range(3):
    ma.train_on_batch(x, y)
    mb.train_on_batch(x.reshape(32, 784), y.reshape(-1, 1))
print("losses", ma.last_step_loss(), mb.last_step_loss())
if "--tc" in sys.argv:
    from tfde_b200.ops import mixed_conv, tensorcore
    a = torch.randn(300, 136, generator=g).to(torch.bfloat16).cuda()
    b = torch.randn(72, 136, generator=g).to(torch.bfloat16).cuda()
    for k in ("tile", "persistent"):
        c = tensorcore.gemm_bf16_tn(a, b, None, True, torch.float32, kernel=k)
    xc = torch.rand(8, 16, 16, 16, generator=g).cuda()
    w = (torch.randn(3, 3, 16, 32, generator=g) * 0.1).cuda()
    yc = mixed_conv.conv2d_fwd(xc, w, None, (1, 1), (1, 1, 1, 1), True)
    mixed_conv.conv2d_bwd(xc, w, yc, torch.ones_like(yc), (1, 1), (1, 1, 1, 1), True, True, False)
    torch.cuda.synchronize()
    print("tc ok", float(c.sum()), float(yc.sum()))
if "--act16" in sys.argv:
    # bf16-activation kernels, MN-major / untransposed-weight GEMMs, staged epilogues
    from tfde_b200 import ops
    from tfde_b200.ops import act16, tensorcore
    BF = torch.bfloat16
    dev = torch.device("cuda")
    xb = torch.randn(8, 16, 16, 64, generator=g).to(BF).to(dev)
    gam, bet = torch.ones(64, device=dev), torch.zeros(64, device=dev)
    mm, mv = torch.zeros(64, device=dev), torch.ones(64, device=dev)
    yb, sm, si, sc, sh = act16.bn_fwd(xb, gam, bet, mm, mv, 0.99, 1e-3, True, True, True)
    act16.bn_bwd(xb, yb, gam, sm, si, sc, sh, True)
    act16.add(xb, yb, relu=True); act16.relu_bwd(yb, xb, addend=xb); act16.colsum(xb.reshape(-1, 64))
    p, idx = act16.maxpool_fwd(xb, (3, 3), (2, 2), (0, 1, 0, 1))
    act16.maxpool_bwd(p, idx, tuple(xb.shape), (3, 3), (2, 2), (0, 1, 0, 1))
    act16.global_avgpool_bwd(act16.global_avgpool_fwd(xb), tuple(xb.shape))
    for (k, s, co) in ((1, 1, 256), (3, 1, 64), (1, 2, 128)):
        ho, wo, pads = ops.conv_geometry(16, 16, k, k, (s, s), "same")
        w = (torch.randn(k, k, 64, co, generator=g) * 0.05).to(dev)
        yc = act16.conv2d_fwd(xb, w, None, (s, s), pads)
        act16.conv2d_bwd(xb, w, torch.ones_like(yc), (s, s), pads)
    a = torch.randn(1000, 136, generator=g).to(BF).to(dev)
    tensorcore.gemm_bf16_nt(a, torch.randn(1000, 72, generator=g).to(BF).to(dev))
    tensorcore.gemm_bf16_nn(a, torch.randn(130, 72, generator=g).to(BF).to(dev), None, True, BF)
    torch.cuda.synchronize()
    print("act16 ok")
torch.cuda.synchronize()
print("SANITIZE SMOKE DONE")
