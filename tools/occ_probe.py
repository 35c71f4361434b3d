import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tfde_b200.ops import native
L = native.load(); torch.zeros(1).cuda()
L.tfde_fused_step_occupancy.argtypes = [ctypes.c_int, ctypes.c_size_t]
for tc in (0, 1):
    print("tc", tc, [(s, L.tfde_fused_step_occupancy(tc, s)) for s in (0, 49152, 70208, 100928, 110000)])
