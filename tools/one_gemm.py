"""One tcgen05 GEMM shape, a few launches (the short program ncu --set full is pointed at).
    python tools/one_gemm.py [M N K] [tile|persistent]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from tfde_b200.ops import tensorcore

M, N, K = (int(v) for v in sys.argv[1:4]) if len(sys.argv) >= 4 else (8192, 8192, 8192)
kern = sys.argv[4] if len(sys.argv) >= 5 else None
g = torch.Generator().manual_seed(0)
a = (torch.randn(M, K, generator=g) * 0.5).to(torch.bfloat16).cuda()
b = (torch.randn(N, K, generator=g) * 0.5).to(torch.bfloat16).cuda()
c = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
for _ in range(4):
    tensorcore.gemm_bf16_tn(a, b, None, False, torch.bfloat16, out=c, kernel=kern)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    tensorcore.gemm_bf16_tn(a, b, None, False, torch.bfloat16, out=c, kernel=kern)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
print(f"{M}x{N}x{K} {kern or tensorcore.pick(M, N, K)}: {ms * 1e3:.1f} us, {2.0 * M * N * K / (ms * 1e-3) / 1e12:.1f} TFLOP/s")
