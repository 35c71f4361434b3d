"""Bring-up aid for umma_tf32.cuh: feed one-hot operands through the self-test kernel and print where they land."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tfde_b200.ops import native
L = native.load()
L.tfde_umma_tf32_selftest.argtypes = [ctypes.c_void_p] * 6 + [ctypes.c_int, ctypes.c_void_p]
L.tfde_umma_tf32_selftest.restype = ctypes.c_int
variant = int(sys.argv[1]) if len(sys.argv) > 1 else 0

def run(act, W, dh):
    D1, D2, D3 = torch.zeros(64, 64).cuda(), torch.zeros(32, 64).cuda(), torch.zeros(64, 32).cuda()
    native.check(L.tfde_umma_tf32_selftest(act.cuda().data_ptr(), W.cuda().data_ptr(), dh.cuda().data_ptr(), D1.data_ptr(),
                                           D2.data_ptr(), D3.data_ptr(), variant, native.stream_ptr()), "selftest")
    torch.cuda.synchronize()
    return D1.cpu(), D2.cpu(), D3.cpu()

# D1[b, n] = sum_c act[b, c] W[c, n]: act one-hot at (b0, c0), W = all ones in row c0 scaled by (n + 1)
for (b0, c0) in [(0, 0), (1, 0), (4, 0), (5, 0), (8, 0), (33, 0), (0, 1), (0, 4), (0, 8), (3, 9), (63, 31)]:
    act = torch.zeros(64, 32); act[b0, c0] = 1.0
    W = torch.zeros(32, 64); W[c0] = torch.arange(1, 65).float()
    D1, _, _ = run(act, W, torch.zeros(64, 64))
    nz = D1.nonzero()
    rows = sorted(set(nz[:, 0].tolist()))
    print(f"act1hot b0={b0} c0={c0}: nonzero rows {rows[:6]} count {len(nz)}; row values {D1[rows[0]][:10].tolist() if rows else None}")
