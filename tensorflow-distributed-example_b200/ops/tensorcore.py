"""Front door of the tcgen05 bf16 GEMMs: picks between the two hand-written kernels.

* ``gemm_tcgen05.cu``            one (128*MT) x BN tile per CTA; MT=2 stacks two accumulators that share
                                 every B tile -- best for large square problems bound by L2->SM operand traffic.
* ``gemm_tcgen05_persistent.cu`` one CTA per SM looping over tiles with a double-buffered TMEM accumulator
                                 (epilogue overlaps the next tile's main loop) -- best for the many
                                 small-K / output-bound GEMMs of a convolutional network.
"""
from __future__ import annotations

import os

import torch

from . import native

c_p, c_i, c_l = native.c_p, native.c_i, native.c_l
native.register_optional("tfde_gemm_bf16_tn_persistent",
                         [c_p, c_p, c_p, c_p, c_l, c_p, c_i, c_i, c_i, c_i, c_p, native.c_z, c_p])

_FORCE = os.environ.get("TFDE_TC_KERNEL", "auto")   # auto | tile | persistent


def pick(M: int, N: int, K: int) -> str:
    if _FORCE in ("tile", "persistent"):
        return _FORCE
    bn = 64 if N <= 64 else (128 if N <= 128 else 256)
    tiles256 = ((M + 255) // 256) * ((N + bn - 1) // bn)
    if bn == 256 and tiles256 >= 2 * 148 and K >= 1024:
        return "tile"          # 256x256 dual-accumulator tiles: a third less operand traffic per FLOP
    return "persistent"


def gemm_bf16_tn(a, b, bias=None, relu=False, out_dtype=torch.float32, out=None, kernel=None):
    """C[M,N] = act(A[M,K] @ B[N,K]^T + bias), bf16 operands (K contiguous, K % 8 == 0), fp32 accumulate."""
    L = native.load()
    if a.dtype != torch.bfloat16 or b.dtype != torch.bfloat16:
        raise TypeError("gemm_bf16_tn expects bf16 operands")
    if not (a.is_contiguous() and b.is_contiguous()):
        raise ValueError("operands must be contiguous")
    M, K = a.shape
    N, K2 = b.shape
    if K != K2:
        raise ValueError(f"K mismatch: {K} vs {K2}")
    c = out if out is not None else torch.empty((M, N), dtype=out_dtype, device=a.device)
    f32 = c.data_ptr() if c.dtype == torch.float32 else None
    b16 = c.data_ptr() if c.dtype == torch.bfloat16 else None
    which = kernel or pick(M, N, K)
    if which == "persistent":
        wsp = native.Workspace.current(a.device)
        native.check(L.tfde_gemm_bf16_tn_persistent(a.data_ptr(), b.data_ptr(), f32, b16, c.stride(0), native.ptr(bias),
                                                    1 if relu else 0, M, N, K, wsp.ws.data_ptr(), wsp.ws_bytes,
                                                    native.stream_ptr()), "gemm_bf16_tn[persistent]")
    else:
        native.check(L.tfde_gemm_bf16_tn(a.data_ptr(), b.data_ptr(), f32, b16, c.stride(0), native.ptr(bias),
                                         1 if relu else 0, M, N, K, native.stream_ptr()), "gemm_bf16_tn[tile]")
    return c


native.register_optional("tfde_gemm_bf16_nt_persistent",
                         [c_p, c_p, c_p, c_p, c_l, c_i, c_i, c_i, c_l, c_l, c_p, native.c_z, c_p])


def gemm_bf16_nt(a, b, out_dtype=torch.float32, out=None):
    """C[M,N] = A[K,M]^T @ B[K,N]: both operands row-major with the REDUCED dimension as rows ("MN-major" for
    tcgen05). This is the filter-gradient GEMM of a convolution / dense layer with its operands exactly as
    the forward pass left them (activations [rows, Cin], output gradients [rows, Cout]) -- no transposed
    copies. M % 8 == 0, N % 8 == 0, any K; row strides may exceed M / N (multiples of 8)."""
    L = native.load()
    if a.dtype != torch.bfloat16 or b.dtype != torch.bfloat16:
        raise TypeError("gemm_bf16_nt expects bf16 operands")
    K, M = a.shape
    K2, N = b.shape
    if K != K2:
        raise ValueError(f"K mismatch: {K} vs {K2}")
    if a.stride(1) != 1 or b.stride(1) != 1:
        raise ValueError("operands must be row-major")
    c = out if out is not None else torch.empty((M, N), dtype=out_dtype, device=a.device)
    f32 = c.data_ptr() if c.dtype == torch.float32 else None
    b16 = c.data_ptr() if c.dtype == torch.bfloat16 else None
    wsp = native.Workspace.current(a.device)
    native.check(L.tfde_gemm_bf16_nt_persistent(a.data_ptr(), b.data_ptr(), f32, b16, c.stride(0), M, N, K,
                                                a.stride(0), b.stride(0), wsp.ws.data_ptr(), wsp.ws_bytes,
                                                native.stream_ptr()), "gemm_bf16_nt[persistent]")
    return c


native.register_optional("tfde_gemm_bf16_nn_persistent",
                         [c_p, c_p, c_p, c_p, c_l, c_p, c_i, c_i, c_i, c_i, c_i, c_l, c_p, native.c_z, c_p])


def gemm_bf16_nn(a, b, bias=None, relu=False, out_dtype=torch.float32, out=None):
    """C[M,N] = act(A[M,K] @ B[Kb,N] + bias): A K-major as usual, B row-major [Kb, N] ("MN-major") -- the weight
    matrix of a convolution / dense layer as Keras stores it, so the forward GEMM needs no transposed weight copy.
    Kb <= K (A may carry zero-padded columns up to a multiple of 8); N % 8 == 0."""
    L = native.load()
    if a.dtype != torch.bfloat16 or b.dtype != torch.bfloat16:
        raise TypeError("gemm_bf16_nn expects bf16 operands")
    M, K = a.shape
    Kb, N = b.shape
    if Kb > K or not a.is_contiguous() or b.stride(1) != 1:
        raise ValueError("gemm_bf16_nn: A must be contiguous [M,K], B row-major [Kb<=K, N]")
    c = out if out is not None else torch.empty((M, N), dtype=out_dtype, device=a.device)
    f32 = c.data_ptr() if c.dtype == torch.float32 else None
    b16 = c.data_ptr() if c.dtype == torch.bfloat16 else None
    wsp = native.Workspace.current(a.device)
    native.check(L.tfde_gemm_bf16_nn_persistent(a.data_ptr(), b.data_ptr(), f32, b16, c.stride(0), native.ptr(bias),
                                                1 if relu else 0, M, N, K, Kb, b.stride(0), wsp.ws.data_ptr(),
                                                wsp.ws_bytes, native.stream_ptr()), "gemm_bf16_nn[persistent]")
    return c


native.register_optional("tfde_gemm_bf16_nn_stats_persistent",
                         [c_p, c_p, c_p, c_l, c_p, c_i, c_i, c_i, c_i, c_l, c_p, c_p, c_p])


def nn_stats_eligible(M: int, N: int) -> bool:
    """The fused-statistics GEMM runs without split-K: worth it only when the output alone fills the machine."""
    bn = 64 if N <= 64 else (128 if N <= 128 else 256)
    return N % 64 == 0 and ((M + 127) // 128) * ((N + bn - 1) // bn) >= 148


def gemm_bf16_nn_stats(a, b, bias, out, stat_sum, stat_sq):
    """gemm_bf16_nn with bf16 output whose epilogue also leaves per-32-row partial column sums / sums of squares
    of the rounded output in ``stat_sum`` / ``stat_sq`` (fp32 [ceil(M/32), N]): the batch-norm that consumes the
    output finalises its statistics from those instead of re-reading the tensor."""
    L = native.load()
    M, K = a.shape
    Kb, N = b.shape
    G = (M + 31) // 32
    if (out.dtype != torch.bfloat16 or stat_sum.shape != (G, N) or stat_sq.shape != (G, N)
            or stat_sum.dtype != torch.float32 or not a.is_contiguous() or b.stride(1) != 1):
        raise ValueError("gemm_bf16_nn_stats: bad operands")
    native.check(L.tfde_gemm_bf16_nn_stats_persistent(a.data_ptr(), b.data_ptr(), out.data_ptr(), out.stride(0),
                                                      native.ptr(bias), M, N, K, Kb, b.stride(0), stat_sum.data_ptr(),
                                                      stat_sq.data_ptr(), native.stream_ptr()),
                 "gemm_bf16_nn_stats[persistent]")
    return out
