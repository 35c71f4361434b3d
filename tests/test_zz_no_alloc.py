"""The allocation-free dense entry point (``tfde_dense_fwd_static_scratch``), exercised the way an
external C caller would: straight through the C ABI with raw device pointers.

Some users of libtfde_kernels.so only own the buffers they are given -- a plug-in running inside a
process whose GPU memory pool belongs to another framework, or a latency-critical loop where
cudaMalloc is forbidden. For them the library carries its own split-K scratch in a static device array.
This test checks that the variant computes the same result as the regular entry point (which gets its
scratch from the caller) on two consecutive calls, that the second call is not disturbed by what the
first one left in the shared array, and that a problem too large for the built-in scratch is refused
instead of being run with too little memory.

Collected last on purpose (file name): a failure here must not hide any other test under ``-x``.
"""
import ctypes

import pytest
import torch

from tfde_b200.ops import native

pytestmark = pytest.mark.gpu

c_p, c_i = ctypes.c_void_p, ctypes.c_int


def _entry_points():
    lib = native.load()
    fn = lib.tfde_dense_fwd_static_scratch
    fn.argtypes = [c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_p]
    fn.restype = c_i
    return lib, fn


def _regular(lib, x, w, b, relu):
    M, K = x.shape
    N = w.shape[1]
    y = torch.empty((M, N), dtype=torch.float32, device=x.device)
    ws = torch.empty(16 << 20, dtype=torch.uint8, device=x.device)
    tickets = torch.zeros(native.MAX_COUNTERS, dtype=torch.int32, device=x.device)
    rc = lib.tfde_dense_fwd(x.data_ptr(), w.data_ptr(), b.data_ptr() if b is not None else None, y.data_ptr(),
                            M, K, N, relu, ws.data_ptr(), ws.numel(), tickets.data_ptr(),
                            torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    return y


def test_static_scratch_matches_caller_scratch_twice_in_a_row():
    lib, fn = _entry_points()
    dev = torch.device("cuda", torch.cuda.current_device())
    g = torch.Generator().manual_seed(0)
    stream = torch.cuda.current_stream().cuda_stream
    # two different split-K problems back to back on one stream: the second must not see the first's leftovers
    for (M, K, N, relu, bias) in [(64, 5408, 64, 1, True), (37, 1176, 200, 0, False), (64, 5408, 64, 1, True)]:
        x = torch.randn(M, K, generator=g).to(dev)
        w = (torch.randn(K, N, generator=g) * 0.05).to(dev)
        b = torch.randn(N, generator=g).to(dev) if bias else None
        y = torch.full((M, N), float("nan"), dtype=torch.float32, device=dev)
        rc = fn(x.data_ptr(), w.data_ptr(), b.data_ptr() if bias else None, y.data_ptr(), M, K, N, relu, stream)
        assert rc == 0, f"entry point returned CUDA error {rc}"
        expect = _regular(lib, x, w, b, relu)
        torch.cuda.synchronize()
        assert torch.isfinite(y).all()
        scale = expect.abs().max().item() + 1e-12
        err = (y - expect).abs().max().item()
        assert err <= 2e-5 + 2e-4 * scale, f"M={M} K={K} N={N}: max abs err {err:.3e} (scale {scale:.3e})"


def test_static_scratch_refuses_problems_beyond_its_capacity():
    lib, fn = _entry_points()
    dev = torch.device("cuda", torch.cuda.current_device())
    M, K, N = 1024, 5408, 64   # split-K schedule wants ~4.5 MiB of partial sums; the array holds 2 MiB
    x = torch.zeros(M, K, device=dev)
    w = torch.zeros(K, N, device=dev)
    y = torch.full((M, N), 7.0, device=dev)
    rc = fn(x.data_ptr(), w.data_ptr(), None, y.data_ptr(), M, K, N, 0, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert rc == 1, "oversized problem must be rejected with cudaErrorInvalidValue"
    assert bool((y == 7.0).all()), "a rejected call must not launch anything"
