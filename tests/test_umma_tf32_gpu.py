"""TF32 tcgen05 GEMMs fed from thread-written shared memory (csrc/kernels/umma_tf32.cuh): the three operand-major
combinations used by the whole-step kernel against an fp32 matmul of TF32-truncated inputs."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _tf32(x):
    return (x.view(torch.int32) & ~0x1FFF).view(torch.float32)


def _run(variant=0, seed=0):
    from tfde_b200.ops import native
    L = native.load()
    L.tfde_umma_tf32_selftest.argtypes = [ctypes.c_void_p] * 6 + [ctypes.c_int, ctypes.c_void_p]
    L.tfde_umma_tf32_selftest.restype = ctypes.c_int
    g = torch.Generator().manual_seed(seed)
    act = torch.randn(64, 32, generator=g).cuda()
    W = torch.randn(32, 64, generator=g).cuda()
    dh = torch.randn(64, 64, generator=g).cuda()
    D1, D2, D3 = torch.zeros(64, 64).cuda(), torch.zeros(32, 64).cuda(), torch.zeros(64, 32).cuda()
    native.check(L.tfde_umma_tf32_selftest(act.data_ptr(), W.data_ptr(), dh.data_ptr(), D1.data_ptr(), D2.data_ptr(),
                                           D3.data_ptr(), variant, native.stream_ptr()), "umma_tf32_selftest")
    torch.cuda.synchronize()
    a, w, d = _tf32(act).double(), _tf32(W).double(), _tf32(dh).double()
    errs = {}
    for name, got, want in (("D1 = act.W", D1, a @ w), ("D2 = act^T.dh", D2, a.t() @ d), ("D3 = dh.W^T", D3, d @ w.t())):
        errs[name] = (got.double() - want).abs().max().item()
    return errs


def test_umma_tf32_three_gemm_forms():
    errs = _run(0)
    print(errs)
    for name, err in errs.items():
        assert err < 2e-3, (name, err)


if __name__ == "__main__":
    for v in range(4):
        print("variant", v, _run(v))
