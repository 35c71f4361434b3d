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
torch.cuda.synchronize()
print("SANITIZE SMOKE DONE")
