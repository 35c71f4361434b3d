"""Mixed-precision dense layer on the tcgen05 tensor-core GEMM (``csrc/kernels/gemm_tcgen05.cu``).

Policy ``mixed_bfloat16`` (BASELINE.json config 3): master weights, optimizer and reductions stay fp32;
the three GEMMs of a Dense layer run on bf16 operands with fp32 accumulation in TMEM. All three map onto
the one K-major ("TN") kernel; fp32->bf16 casts (with the ReLU-backward mask folded in) and transposes
are produced by small glue kernels (``mixed_ops.cu``).

The CPU restatement rounds the same operands to bf16 and multiplies in fp32 -- that is the oracle the GPU
path is tested against.
"""
from __future__ import annotations

import torch

from . import native
from . import reference as ref
from . import tensorcore

c_p, c_i, c_l = native.c_p, native.c_i, native.c_l
native.register_optional("tfde_cast_bf16_masked", [c_p, c_p, c_p, c_l, c_p])
native.register_optional("tfde_colsum_masked", [c_p, c_p, c_p, c_l, c_i, c_p, c_p, c_p])

_policy = "float32"
_activations = "float32"


def set_global_policy(name: str, activations: str = None):
    """tf.keras.mixed_precision.set_global_policy: 'float32' (reference default) or 'mixed_bfloat16'.

    ``activations`` (extension): dtype of the tensors handed from layer to layer under 'mixed_bfloat16'.
    'float32' (default) keeps fp32 activations and rounds only the GEMM operands; 'bfloat16' is Keras' own
    mixed-precision semantics (bf16 layer outputs, fp32 variables) and what convolutional networks should
    use: see ops/act16.py."""
    global _policy, _activations
    if name not in ("float32", "mixed_bfloat16"):
        raise ValueError("policy must be 'float32' or 'mixed_bfloat16'")
    if activations not in (None, "float32", "bfloat16"):
        raise ValueError("activations must be 'float32' or 'bfloat16'")
    if activations == "bfloat16" and name != "mixed_bfloat16":
        raise ValueError("bfloat16 activations need the 'mixed_bfloat16' policy")
    _policy = name
    _activations = activations or "float32"


def global_policy() -> str:
    return _policy


def activation_dtype() -> str:
    return _activations if _policy == "mixed_bfloat16" else "float32"


def supported(M: int, K: int, N: int) -> bool:
    """All three GEMMs need their reduction dimension to be a multiple of 8 (TMA 16-byte strides); below
    ~128 rows the layer is latency-bound and stays on the fp32 path."""
    return M >= 128 and M % 8 == 0 and K % 8 == 0 and N % 8 == 0


def _bf(x):
    return x.to(torch.bfloat16).to(torch.float32)


def cast_bf16_masked(x, mask=None, out=None):
    L = native.load()
    y = out if out is not None else torch.empty(x.shape, dtype=torch.bfloat16, device=x.device)
    native.check(L.tfde_cast_bf16_masked(x.data_ptr(), native.ptr(mask), y.data_ptr(), x.numel(),
                                         native.stream_ptr()), "cast_bf16_masked")
    return y


def colsum_masked(x, mask=None, out=None):
    L = native.load()
    R, C = x.shape
    o = out if out is not None else torch.empty((C,), dtype=torch.float32, device=x.device)
    wsp = native.Workspace.current(x.device)
    native.check(L.tfde_colsum_masked(x.data_ptr(), native.ptr(mask), o.data_ptr(), R, C, wsp.small.data_ptr(),
                                      wsp.counters.data_ptr(), native.stream_ptr()), "colsum_masked")
    return o


def dense_fwd(x, w, bias=None, act="none", out=None, bufs=None):
    """y = act(bf16(x) @ bf16(w) + bias), fp32 result."""
    relu = act == "relu"
    if act not in ("none", "relu"):
        raise ValueError("mixed dense supports 'none' and 'relu'")
    if not x.is_cuda:
        y = _bf(x) @ _bf(w)
        if bias is not None:
            y = y + bias
        y = torch.relu(y) if relu else y
        if out is not None:
            out.copy_(y)
            return out
        return y
    b = bufs if bufs is not None else {}
    xb = cast_bf16_masked(x, None, out=b.get("xb"))
    wt = native.cast_transpose_bf16(w, None, out=b.get("wt"))        # [N, K]
    return tensorcore.gemm_bf16_tn(xb, wt, bias, relu, torch.float32, out=out)


def dense_bwd(x, w, y, dy, act="none", need_dx=True, has_bias=True, out_dx=None, out_dw=None, out_db=None,
              bufs=None):
    """Returns (dx|None, dw, db|None); ReLU backward is folded into the casts of dy."""
    relu = act == "relu"
    if not x.is_cuda:
        g = dy * (y > 0).to(dy.dtype) if relu else dy
        dw = _bf(x).t() @ _bf(g)
        db = g.sum(0) if has_bias else None
        dx = _bf(g) @ _bf(w).t() if need_dx else None
        for o, v in ((out_dx, dx), (out_dw, dw), (out_db, db)):
            if o is not None and v is not None:
                o.copy_(v)
        return (out_dx if out_dx is not None else dx, out_dw if out_dw is not None else dw,
                (out_db if out_db is not None else db))
    b = bufs if bufs is not None else {}
    mask = y if relu else None
    M, K = x.shape
    N = w.shape[1]
    xt = native.cast_transpose_bf16(x, None, out=b.get("xt"))          # [K, M]
    dyt = native.cast_transpose_bf16(dy, mask, out=b.get("dyt"))       # [N, M]
    dw = tensorcore.gemm_bf16_tn(xt, dyt, None, False, torch.float32, out=out_dw)   # [K, N]
    db = colsum_masked(dy, mask, out=out_db) if has_bias else None
    dx = None
    if need_dx:
        dyb = cast_bf16_masked(dy, mask, out=b.get("dyb"))             # [M, N]
        wb = cast_bf16_masked(w, None, out=b.get("wb"))                # [K, N]
        dx = tensorcore.gemm_bf16_tn(dyb, wb, None, False, torch.float32, out=out_dx)   # [M, K]
    return dx, dw, db
