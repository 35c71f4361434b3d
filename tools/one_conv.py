"""One implicit-GEMM convolution shape: forward, data gradient, filter gradient, a few launches each with device
timing (the short program ncu --set full is pointed at).
    python tools/one_conv.py [B H W C Cout [k]]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from tfde_b200.ops import conv_igemm

B, H, W, C, CO = (int(v) for v in sys.argv[1:6]) if len(sys.argv) >= 6 else (64, 56, 56, 64, 64)
k = int(sys.argv[6]) if len(sys.argv) >= 7 else 3
p = k // 2
pads, stride = (p, p, p, p), (1, 1)
g = torch.Generator().manual_seed(0)
x = torch.randn(B, H, W, C, generator=g).to(torch.bfloat16).cuda()
w = (torch.randn(k, k, C, CO, generator=g) * 0.05).to(torch.bfloat16).cuda()
dy = torch.randn(B, H, W, CO, generator=g).to(torch.bfloat16).cuda()
y = torch.empty(B, H, W, CO, dtype=torch.bfloat16, device="cuda")
dx = torch.empty_like(x)
dw = torch.empty(k, k, C, CO, dtype=torch.float32, device="cuda")
ops = {"fwd": lambda: conv_igemm.fwd(x, w, None, stride, pads, y),
       "dgrad": lambda: conv_igemm.dgrad(dy, w, pads, dx),
       "wgrad": lambda: conv_igemm.wgrad(x, dy, tuple(w.shape), stride, pads, dw)}
for _ in range(2):
    for f in ops.values():
        f()
torch.cuda.synchronize()
flop = 2.0 * B * H * W * k * k * C * CO
io = {"fwd": (x.numel() + y.numel()) * 2, "dgrad": (dy.numel() + dx.numel()) * 2, "wgrad": (x.numel() + dy.numel()) * 2}
for name, f in ops.items():
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        f()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / 5 * 1e3
    print(f"{name:5s} B{B} {H}x{W} {C}->{CO} k{k}: {us:7.1f} us  {flop / us / 1e6:7.1f} TFLOP/s  "
          f"compulsory HBM traffic {io[name] / 1e6:.1f} MB = {io[name] / us / 1e3:.0f} GB/s")
