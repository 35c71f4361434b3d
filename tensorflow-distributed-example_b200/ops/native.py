"""ctypes binding of ``libtfde_kernels.so`` (the hand-written sm_100a kernels in ``csrc/kernels``).

Every wrapper takes/returns torch CUDA tensors, launches on the *current* torch stream (so the
launches are captured by ``torch.cuda.graph``), and mirrors the signature of the same-named function
in :mod:`.reference`. There is deliberately no fallback here: if the library is missing on a machine
with a GPU the import of a CUDA op raises (a silent eager fallback would hide a broken build).

Scratch memory (split-K partials, reduction partials, self-cleaning ticket counters) comes from a
per-(device, stream) :class:`Workspace` so that kernels on different streams never share tickets.
"""
from __future__ import annotations

import ctypes
import os
import threading
from typing import Optional, Tuple

import torch

_LIB_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "_lib")
_LIB_PATH = os.path.join(_LIB_DIR, "libtfde_kernels.so")
_lib = None
_lock = threading.Lock()

c_p = ctypes.c_void_p
c_i = ctypes.c_int
c_l = ctypes.c_long
c_f = ctypes.c_float
c_z = ctypes.c_size_t
c_u = ctypes.c_uint

MAX_COUNTERS = 4096
HPARAMS_FLOATS = 16
# indices into the hparams block (csrc/kernels/common.cuh::TfdeHParams)
HP_LR, HP_MOMENTUM, HP_BETA1, HP_BETA2, HP_EPS, HP_GRAD_SCALE, HP_B1POW, HP_B2POW = range(8)
HP_STEP, HP_SEED_LO, HP_SEED_HI = 8, 9, 10

OPT_SGD, OPT_MOMENTUM, OPT_NESTEROV, OPT_ADAM = 0, 1, 2, 3


class NativeLibraryError(RuntimeError):
    pass


def lib_path() -> str:
    return _LIB_PATH


def available() -> bool:
    return os.path.exists(_LIB_PATH)


def load():
    """Loads (once) and returns the ctypes handle. Raises NativeLibraryError if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(_LIB_PATH):
            raise NativeLibraryError(
                f"{_LIB_PATH} is missing: run `python build.py` (or __graft_entry__.build()) first. "
                "CUDA tensors are only served by the native sm_100a kernels; there is no eager fallback.")
        lib = ctypes.CDLL(_LIB_PATH)
        _declare(lib)
        _lib = lib
    return _lib


def _declare(lib):
    def sig(name, args, res=c_i):
        fn = getattr(lib, name)
        fn.argtypes = args
        fn.restype = res

    gemm_tail = [c_p, c_z, c_p, c_p]  # ws, ws_bytes, counters, stream
    sig("tfde_dense_fwd", [c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_i] + gemm_tail)
    sig("tfde_dense_wgrad", [c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_i] + gemm_tail)
    sig("tfde_dense_dgrad", [c_p, c_p, c_p, c_p, c_i, c_i, c_i] + gemm_tail)
    geom = [c_i] * 13  # Bn,H,W,C,KH,KW,Ho,Wo,sh,sw,pt,pl,Cout
    sig("tfde_conv2d_fwd", [c_p, c_p, c_p, c_p] + geom + [c_i] + gemm_tail)
    sig("tfde_conv2d_wgrad", [c_p, c_p, c_p, c_p, c_p] + geom + gemm_tail)
    sig("tfde_conv2d_dgrad", [c_p, c_p, c_p, c_p] + geom + gemm_tail)
    sig("tfde_gemm_max_counters", [], c_i)
    sig("tfde_conv3x3c1_relu_pool_fwd", [c_p, c_i, c_f, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_p])
    sig("tfde_conv3x3c1_relu_pool_bwd_ws_floats", [], c_z)
    sig("tfde_conv3x3c1_relu_pool_bwd",
        [c_p, c_i, c_f, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_p])
    sig("tfde_bn_ws_floats", [c_l, c_i], c_z)
    sig("tfde_bn_fwd_train", [c_p, c_p, c_l, c_i, c_p, c_p, c_p, c_p, c_p, c_p, c_f, c_f, c_i, c_i,
                              c_p, c_p, c_p])
    sig("tfde_bn_fwd_infer", [c_p, c_p, c_l, c_i, c_p, c_p, c_p, c_p, c_f, c_i, c_p])
    sig("tfde_bn_bwd", [c_p, c_p, c_p, c_p, c_l, c_i, c_p, c_p, c_p, c_i, c_p, c_p, c_p, c_p, c_p])
    pool = [c_i] * 13  # Bn,H,W,C,PH,PW,sh,sw,pt,pl,Ho,Wo  (12) + stream below
    sig("tfde_maxpool_fwd", [c_p, c_p, c_p] + [c_i] * 12 + [c_p])
    sig("tfde_maxpool_bwd", [c_p, c_p, c_p] + [c_i] * 12 + [c_p])
    sig("tfde_relu_fwd", [c_p, c_p, c_l, c_p])
    sig("tfde_relu_bwd", [c_p, c_p, c_p, c_l, c_p])
    sig("tfde_add", [c_p, c_p, c_p, c_l, c_i, c_p])
    sig("tfde_u8_to_f32", [c_p, c_p, c_l, c_f, c_p])
    sig("tfde_dropout", [c_p, c_p, c_l, c_f, c_p, c_u, c_i, c_p])
    sig("tfde_softmax_fwd", [c_p, c_p, c_i, c_i, c_p])
    sig("tfde_softmax_bwd", [c_p, c_p, c_p, c_i, c_i, c_p])
    sig("tfde_gap_fwd", [c_p, c_p, c_i, c_i, c_i, c_p])
    sig("tfde_gap_bwd", [c_p, c_p, c_i, c_i, c_i, c_p])
    sig("tfde_xent_ws_floats", [], c_z)
    sig("tfde_xent", [c_p, c_p, c_i, c_p, c_i, c_i, c_f, c_i, c_p, c_p, c_p, c_p])
    sig("tfde_optimizer_apply", [c_p, c_p, c_p, c_p, c_p, c_l, c_i, c_p, c_p, c_i, c_p, c_p, c_p])
    sig("tfde_hparams_advance", [c_p, c_p])
    for name, args, res in _OPTIONAL_SIGS:
        if hasattr(lib, name):
            sig(name, args, res)


# Entry points that live in later-built translation units (declared if present).
_OPTIONAL_SIGS = []


def register_optional(name, args, res=c_i):
    _OPTIONAL_SIGS.append((name, args, res))
    if _lib is not None and hasattr(_lib, name):
        fn = getattr(_lib, name)
        fn.argtypes = args
        fn.restype = res


_launches = 0
_trace_sink = None


def set_trace_sink(sink):
    """While set, every native entry point that launches records its name into ``sink`` (obs.profiler.StepTracer).
    Returns the previous sink."""
    global _trace_sink
    prev, _trace_sink = _trace_sink, sink
    return prev


def check(code: int, what: str, kernels: int = 1):
    """Raises on a CUDA error; counts the kernels this entry point launched (bench.py reports them)."""
    global _launches
    if code != 0:
        raise RuntimeError(f"{what} failed with CUDA error {code}")
    _launches += kernels
    if _trace_sink is not None and kernels:
        _trace_sink.append("tfde_" + what)


def launch_count() -> int:
    return _launches


def register_wgrad_push(grad_ptr: int, desc) -> None:
    from . import push as _push
    _push.register(grad_ptr, desc)


def stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _f32c(t: torch.Tensor, name: str) -> torch.Tensor:
    if t.dtype != torch.float32:
        raise TypeError(f"{name}: expected float32, got {t.dtype}")
    if not t.is_contiguous():
        raise ValueError(f"{name}: tensor must be contiguous (NHWC / row-major)")
    return t


class Workspace:
    """Scratch for one (device, stream): split-K partials + reduction partials + ticket counters."""

    _cache = {}

    def __init__(self, device, ws_bytes=64 << 20):
        self.ws = torch.empty(ws_bytes // 4, dtype=torch.float32, device=device)
        self.ws_bytes = self.ws.numel() * 4
        self.small = torch.empty(1 << 18, dtype=torch.float32, device=device)  # 1 MB of partials
        self.counters = torch.zeros(MAX_COUNTERS, dtype=torch.int32, device=device)

    @classmethod
    def current(cls, device=None) -> "Workspace":
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        key = (dev.index if dev.index is not None else torch.cuda.current_device(),
               torch.cuda.current_stream(dev).cuda_stream)
        w = cls._cache.get(key)
        if w is None:
            # allocate outside of any ongoing graph capture
            w = cls(dev)
            cls._cache[key] = w
        return w

    @classmethod
    def preallocate(cls, device=None):
        return cls.current(device)


# ----------------------------------------------------------------------------------------------
# hparams block
# ----------------------------------------------------------------------------------------------
def make_hparams(device, lr=0.01, momentum=0.0, beta1=0.9, beta2=0.999, eps=1e-7, grad_scale=1.0,
                 seed=0, step=0) -> torch.Tensor:
    """Device block read by the optimizer / dropout kernels (TfdeHParams). Returned as 16 x int32 view
    pair: use :func:`hparams_set` to update fields without re-capturing a CUDA graph."""
    import struct
    raw = struct.pack("<8f8I", lr, momentum, beta1, beta2, eps, grad_scale,
                      beta1 ** step, beta2 ** step, step, seed & 0xFFFFFFFF,
                      (seed >> 32) & 0xFFFFFFFF, 0, 0, 0, 0, 0)
    host = torch.frombuffer(bytearray(raw), dtype=torch.int32).clone()
    return host.to(device)


def hparams_set(hp: torch.Tensor, **fields):
    """Asynchronous field update (stream ordered)."""
    import struct
    names = {"lr": HP_LR, "momentum": HP_MOMENTUM, "beta1": HP_BETA1, "beta2": HP_BETA2,
             "eps": HP_EPS, "grad_scale": HP_GRAD_SCALE}
    for k, v in fields.items():
        if k in names:
            bits = struct.unpack("<i", struct.pack("<f", float(v)))[0]
            hp[names[k]].fill_(bits)
        elif k == "step":
            hp[HP_STEP].fill_(int(v))
        else:
            raise KeyError(k)


def hparams_get(hp: torch.Tensor) -> dict:
    import struct
    raw = hp.cpu().numpy().tobytes()
    f = struct.unpack("<8f8I", raw)
    return dict(lr=f[0], momentum=f[1], beta1=f[2], beta2=f[3], eps=f[4], grad_scale=f[5],
                beta1_pow=f[6], beta2_pow=f[7], step=f[8], seed=f[9] | (f[10] << 32))


# ----------------------------------------------------------------------------------------------
# GEMM family
# ----------------------------------------------------------------------------------------------
def _gemm_tail(wsp: Workspace):
    return (wsp.ws.data_ptr(), wsp.ws_bytes, wsp.counters.data_ptr(), stream_ptr())


def dense_fwd(x, w, bias=None, act="none", out=None):
    L = load()
    _f32c(x, "x"); _f32c(w, "w")
    M, K = x.shape
    N = w.shape[1]
    y = out if out is not None else torch.empty((M, N), dtype=torch.float32, device=x.device)
    wsp = Workspace.current(x.device)
    if _SKINNY and L.tfde_dense_skinny_supported(M, K, N):
        # small M*N, long K (Dense(64) over 5408 inputs at batch 64): cluster/DSMEM split-K kernel
        check(L.tfde_dense_skinny_fwd(x.data_ptr(), w.data_ptr(), ptr(bias), y.data_ptr(), M, K, N,
                                      1 if act == "relu" else 0, wsp.ws.data_ptr(), wsp.counters.data_ptr(),
                                      stream_ptr()), "dense_skinny_fwd")
    else:
        check(L.tfde_dense_fwd(x.data_ptr(), w.data_ptr(), ptr(bias), y.data_ptr(), M, K, N,
                               1 if act == "relu" else 0, *_gemm_tail(wsp)), "dense_fwd")
    if act == "softmax":
        check(L.tfde_softmax_fwd(y.data_ptr(), y.data_ptr(), M, N, stream_ptr()), "softmax_fwd")
    elif act not in ("none", "relu"):
        raise ValueError(act)
    return y


_SKINNY_BWD = os.environ.get("TFDE_SKINNY_BWD", "1") == "1"
register_optional("tfde_dense_bwd_skinny_supported", [c_i, c_i, c_i])
register_optional("tfde_dense_bwd_skinny", [c_p] * 7 + [c_i] * 3 + [c_p])


def dense_bwd(x, w, y, dy, act="none", need_dx=True, has_bias=True, out_dx=None, out_dw=None,
              out_db=None, scratch=None, side_stream=None):
    """``side_stream``: when given (and a data gradient is needed) the filter-gradient GEMM is forked onto
    it and runs concurrently with the data-gradient GEMM; the caller joins the side stream before anything
    consumes dw/db (the optimizer / all-reduce). Captured CUDA graphs keep the two branches parallel."""
    L = load()
    _f32c(x, "x"); _f32c(w, "w"); _f32c(dy, "dy")
    M, K = x.shape
    N = w.shape[1]
    dev = x.device
    wsp = Workspace.current(dev)
    mask = None
    if act == "relu":
        mask = y
    elif act == "softmax":
        dz = scratch if scratch is not None else torch.empty_like(dy)
        check(L.tfde_softmax_bwd(y.data_ptr(), dy.data_ptr(), dz.data_ptr(), M, N, stream_ptr()),
              "softmax_bwd")
        dy = dz
    dw = out_dw if out_dw is not None else torch.empty((K, N), dtype=torch.float32, device=dev)
    db = None
    if has_bias:
        db = out_db if out_db is not None else torch.empty((N,), dtype=torch.float32, device=dev)
    from . import push as _push
    pd = _push.lookup(dw.data_ptr())
    if (pd is None and _SKINNY_BWD and hasattr(L, "tfde_dense_bwd_skinny")
            and L.tfde_dense_bwd_skinny_supported(M, K, N)):
        # batch <= 64, 64 units: dx, dW and db in ONE launch (csrc/kernels/small_batch.cu) instead of two forked
        # 85-block GEMMs -- the data gradient is on the critical path of the step
        dx = (out_dx if out_dx is not None else torch.empty((M, K), dtype=torch.float32, device=dev)) if need_dx else None
        check(L.tfde_dense_bwd_skinny(x.data_ptr(), w.data_ptr(), dy.data_ptr(), ptr(mask), ptr(dx), dw.data_ptr(),
                                      ptr(db), M, K, N, stream_ptr()), "dense_bwd_skinny")
        return dx, dw, db

    def launch_wgrad(ws_):
        if pd is not None:
            # GEMM -> reduce-scatter in one kernel: tiles also go straight into the peers' receive slots
            check(L.tfde_dense_wgrad_push(x.data_ptr(), dy.data_ptr(), ptr(mask), dw.data_ptr(), ptr(db), M, K, N,
                                          ws_.ws.data_ptr(), ws_.ws_bytes, ws_.counters.data_ptr(), ctypes.byref(pd),
                                          stream_ptr()), "dense_wgrad_push")
            _push.stats["pushed_launches"] += 1
        else:
            check(L.tfde_dense_wgrad(x.data_ptr(), dy.data_ptr(), ptr(mask), dw.data_ptr(), ptr(db), M, K, N,
                                     *_gemm_tail(ws_)), "dense_wgrad")

    if callable(side_stream):      # lazily created: the one-launch path above never needs it
        side_stream = side_stream() if need_dx else None
    if side_stream is not None and need_dx:
        side_stream.wait_stream(torch.cuda.current_stream(dev))      # fork: inputs are ready
        with torch.cuda.stream(side_stream):
            launch_wgrad(Workspace.current(dev))                     # the side stream has its own scratch
    else:
        launch_wgrad(wsp)
    dx = None
    if need_dx:
        dx = out_dx if out_dx is not None else torch.empty((M, K), dtype=torch.float32, device=dev)
        check(L.tfde_dense_dgrad(dy.data_ptr(), ptr(mask), w.data_ptr(), dx.data_ptr(), M, K, N,
                                 *_gemm_tail(wsp)), "dense_dgrad")
    return dx, dw, db


def _geom(x_shape, w_shape, stride, pads, ho, wo):
    bn, h, wd, c = x_shape
    kh, kw, _, cout = w_shape
    return (bn, h, wd, c, kh, kw, ho, wo, stride[0], stride[1], pads[0], pads[2], cout)


def _conv_out(x_shape, w_shape, stride, pads):
    _, h, wd, _ = x_shape
    kh, kw = w_shape[0], w_shape[1]
    ho = (h + pads[0] + pads[1] - kh) // stride[0] + 1
    wo = (wd + pads[2] + pads[3] - kw) // stride[1] + 1
    return ho, wo


register_optional("tfde_smallconv_supported", [c_i] * 15)
register_optional("tfde_smallconv_bwd_ws_floats", [c_i] * 5, c_z)
register_optional("tfde_smallconv_fwd", [c_p, c_p, c_p, c_p] + [c_i] * 14 + [c_p])
register_optional("tfde_smallconv_bwd", [c_p] * 6 + [c_i] * 13 + [c_p])
_SMALLCONV = os.environ.get("TFDE_SMALLCONV", "1") == "1"


def conv2d_fwd(x, w, bias=None, stride=(1, 1), pads=(0, 0, 0, 0), relu=False, out=None):
    L = load()
    _f32c(x, "x"); _f32c(w, "w")
    ho, wo = _conv_out(x.shape, w.shape, stride, pads)
    y = out if out is not None else torch.empty((x.shape[0], ho, wo, w.shape[3]), dtype=torch.float32,
                                                device=x.device)
    g = _geom(x.shape, w.shape, stride, pads, ho, wo)
    if _SMALLCONV and L.tfde_smallconv_supported(*g, 0, 0):
        # small channel counts (Model B's 6/12/24-channel layers): direct convolution, one CTA per image
        check(L.tfde_smallconv_fwd(x.data_ptr(), w.data_ptr(), ptr(bias), y.data_ptr(), *g, 1 if relu else 0,
                                   stream_ptr()), "smallconv_fwd")
        return y
    wsp = Workspace.current(x.device)
    check(L.tfde_conv2d_fwd(x.data_ptr(), w.data_ptr(), ptr(bias), y.data_ptr(),
                            *_geom(x.shape, w.shape, stride, pads, ho, wo), 1 if relu else 0,
                            *_gemm_tail(wsp)), "conv2d_fwd")
    return y


def conv2d_bwd(x, w, y, dy, stride=(1, 1), pads=(0, 0, 0, 0), relu=False, need_dx=True, has_bias=True,
               out_dx=None, out_dw=None, out_db=None):
    L = load()
    _f32c(x, "x"); _f32c(w, "w"); _f32c(dy, "dy")
    dev = x.device
    ho, wo = dy.shape[1], dy.shape[2]
    g = _geom(x.shape, w.shape, stride, pads, ho, wo)
    wsp = Workspace.current(dev)
    mask = y if relu else None
    dw = out_dw if out_dw is not None else torch.empty(tuple(w.shape), dtype=torch.float32, device=dev)
    if (_SMALLCONV and not relu and not has_bias and L.tfde_smallconv_supported(*g, 1, 1 if need_dx else 0)
            and L.tfde_smallconv_bwd_ws_floats(g[0], g[3], g[4], g[5], g[12]) * 4 <= wsp.ws_bytes):
        dx = None
        if need_dx:
            dx = out_dx if out_dx is not None else torch.empty(tuple(x.shape), dtype=torch.float32, device=dev)
        check(L.tfde_smallconv_bwd(x.data_ptr(), dy.data_ptr(), w.data_ptr(), ptr(dx), dw.data_ptr(), wsp.ws.data_ptr(),
                                   *g, stream_ptr()), "smallconv_bwd", kernels=3 if need_dx else 2)
        return dx, dw, None
    db = None
    if has_bias:
        db = out_db if out_db is not None else torch.empty((w.shape[3],), dtype=torch.float32, device=dev)
    check(L.tfde_conv2d_wgrad(x.data_ptr(), dy.data_ptr(), ptr(mask), dw.data_ptr(), ptr(db), *g,
                              *_gemm_tail(wsp)), "conv2d_wgrad")
    dx = None
    if need_dx:
        dx = out_dx if out_dx is not None else torch.empty(tuple(x.shape), dtype=torch.float32, device=dev)
        check(L.tfde_conv2d_dgrad(dy.data_ptr(), ptr(mask), w.data_ptr(), dx.data_ptr(), *g,
                                  *_gemm_tail(wsp)), "conv2d_dgrad")
    return dx, dw, db


# ----------------------------------------------------------------------------------------------
# fused Model-A first layer
# ----------------------------------------------------------------------------------------------
def conv_relu_pool_supported(x, w) -> bool:
    return (x.is_cuda and x.dim() == 4 and x.shape[3] == 1 and tuple(w.shape[:3]) == (3, 3, 1)
            and w.shape[3] == 32 and x.dtype in (torch.float32, torch.uint8))


def conv_relu_pool_fwd(x, w, bias, in_scale=1.0 / 255.0, out=None, out_idx=None):
    L = load()
    if not x.is_contiguous():
        raise ValueError("x must be contiguous")
    bn, h, wd, _ = x.shape
    ph, pw = (h - 2) // 2, (wd - 2) // 2
    pooled = out if out is not None else torch.empty((bn, ph, pw, 32), dtype=torch.float32, device=x.device)
    idx = out_idx if out_idx is not None else torch.empty((bn, ph, pw, 32), dtype=torch.uint8, device=x.device)
    check(L.tfde_conv3x3c1_relu_pool_fwd(x.data_ptr(), 1 if x.dtype == torch.uint8 else 0, in_scale,
                                         w.data_ptr(), ptr(bias), pooled.data_ptr(), idx.data_ptr(),
                                         bn, h, wd, 32, stream_ptr()), "conv_relu_pool_fwd")
    return pooled, idx


def conv_relu_pool_bwd(x, w, pooled, idx, dpooled, need_dx=False, in_scale=1.0 / 255.0, out_dw=None,
                       out_db=None):
    if need_dx:
        raise NotImplementedError("fused conv+relu+pool backward is a first-layer kernel (no dx)")
    L = load()
    bn, h, wd, _ = x.shape
    dev = x.device
    wsp = Workspace.current(dev)
    dw = out_dw if out_dw is not None else torch.empty((3, 3, 1, 32), dtype=torch.float32, device=dev)
    db = out_db if out_db is not None else torch.empty((32,), dtype=torch.float32, device=dev)
    check(L.tfde_conv3x3c1_relu_pool_bwd(x.data_ptr(), 1 if x.dtype == torch.uint8 else 0, in_scale,
                                         pooled.data_ptr(), idx.data_ptr(), dpooled.data_ptr(),
                                         dw.data_ptr(), db.data_ptr(), wsp.small.data_ptr(),
                                         wsp.counters.data_ptr(), bn, h, wd, 32, stream_ptr()),
          "conv_relu_pool_bwd")
    return None, dw, db


# ----------------------------------------------------------------------------------------------
# pooling / norm / elementwise
# ----------------------------------------------------------------------------------------------
def maxpool_fwd(x, pool=(2, 2), stride=(2, 2), pads=(0, 0, 0, 0), out=None, out_idx=None):
    L = load()
    _f32c(x, "x")
    bn, h, wd, c = x.shape
    ho = (h + pads[0] + pads[1] - pool[0]) // stride[0] + 1
    wo = (wd + pads[2] + pads[3] - pool[1]) // stride[1] + 1
    y = out if out is not None else torch.empty((bn, ho, wo, c), dtype=torch.float32, device=x.device)
    idx = out_idx if out_idx is not None else torch.empty((bn, ho, wo, c), dtype=torch.uint8, device=x.device)
    check(L.tfde_maxpool_fwd(x.data_ptr(), y.data_ptr(), idx.data_ptr(), bn, h, wd, c, pool[0], pool[1],
                             stride[0], stride[1], pads[0], pads[2], ho, wo, stream_ptr()), "maxpool_fwd")
    return y, idx


def maxpool_bwd(dy, idx, x_shape, pool=(2, 2), stride=(2, 2), pads=(0, 0, 0, 0), out=None):
    L = load()
    _f32c(dy, "dy")
    bn, h, wd, c = x_shape
    ho, wo = dy.shape[1], dy.shape[2]
    dx = out if out is not None else torch.empty(tuple(x_shape), dtype=torch.float32, device=dy.device)
    check(L.tfde_maxpool_bwd(dy.data_ptr(), idx.data_ptr(), dx.data_ptr(), bn, h, wd, c, pool[0], pool[1],
                             stride[0], stride[1], pads[0], pads[2], ho, wo, stream_ptr()), "maxpool_bwd")
    return dx


def bn_fwd(x, gamma, beta, moving_mean, moving_var, momentum=0.99, eps=1e-3, training=True, relu=False,
           unbiased_moving=True, out=None, save_mean=None, save_invstd=None):
    """In-place update of moving statistics (returns them for signature parity with the reference)."""
    L = load()
    _f32c(x, "x")
    c = x.shape[-1]
    m = x.numel() // c
    dev = x.device
    y = out if out is not None else torch.empty_like(x)
    if training:
        wsp = Workspace.current(dev)
        need = L.tfde_bn_ws_floats(m, c)
        if need > wsp.small.numel():
            raise RuntimeError("bn workspace too small")
        sm = save_mean if save_mean is not None else torch.empty((c,), dtype=torch.float32, device=dev)
        si = save_invstd if save_invstd is not None else torch.empty((c,), dtype=torch.float32, device=dev)
        check(L.tfde_bn_fwd_train(x.data_ptr(), y.data_ptr(), m, c, ptr(gamma), ptr(beta),
                                  moving_mean.data_ptr(), moving_var.data_ptr(), sm.data_ptr(),
                                  si.data_ptr(), momentum, eps, 1 if relu else 0,
                                  1 if unbiased_moving else 0, wsp.small.data_ptr(),
                                  wsp.counters.data_ptr(), stream_ptr()), "bn_fwd_train", kernels=2)
        return y, sm, si, moving_mean, moving_var
    check(L.tfde_bn_fwd_infer(x.data_ptr(), y.data_ptr(), m, c, ptr(gamma), ptr(beta),
                              moving_mean.data_ptr(), moving_var.data_ptr(), eps, 1 if relu else 0,
                              stream_ptr()), "bn_fwd_infer")
    return y, moving_mean, None, moving_mean, moving_var


def bn_bwd(x, y, dy, gamma, save_mean, save_invstd, relu=False, out_dx=None, out_dgamma=None,
           out_dbeta=None):
    L = load()
    _f32c(x, "x"); _f32c(dy, "dy")
    c = x.shape[-1]
    m = x.numel() // c
    dev = x.device
    wsp = Workspace.current(dev)
    dx = out_dx if out_dx is not None else torch.empty_like(x)
    dbeta = out_dbeta if out_dbeta is not None else torch.empty((c,), dtype=torch.float32, device=dev)
    dgamma = out_dgamma if out_dgamma is not None else torch.empty((c,), dtype=torch.float32, device=dev)
    check(L.tfde_bn_bwd(x.data_ptr(), y.data_ptr(), dy.data_ptr(), dx.data_ptr(), m, c, ptr(gamma),
                        save_mean.data_ptr(), save_invstd.data_ptr(), 1 if relu else 0, dbeta.data_ptr(),
                        dgamma.data_ptr(), wsp.small.data_ptr(), wsp.counters.data_ptr(), stream_ptr()),
          "bn_bwd", kernels=2)
    return dx, (dgamma if gamma is not None else None), dbeta


def relu_fwd(x, out=None):
    L = load()
    y = out if out is not None else torch.empty_like(x)
    check(L.tfde_relu_fwd(_f32c(x, "x").data_ptr(), y.data_ptr(), x.numel(), stream_ptr()), "relu_fwd")
    return y


def relu_bwd(y, dy, out=None):
    L = load()
    dx = out if out is not None else torch.empty_like(dy)
    check(L.tfde_relu_bwd(y.data_ptr(), _f32c(dy, "dy").data_ptr(), dx.data_ptr(), dy.numel(), stream_ptr()),
          "relu_bwd")
    return dx


def softmax_fwd(z, out=None):
    L = load()
    p = out if out is not None else torch.empty_like(z)
    check(L.tfde_softmax_fwd(_f32c(z, "z").data_ptr(), p.data_ptr(), z.shape[0], z.shape[1], stream_ptr()),
          "softmax_fwd")
    return p


def softmax_bwd(p, dp, out=None):
    L = load()
    dz = out if out is not None else torch.empty_like(dp)
    check(L.tfde_softmax_bwd(p.data_ptr(), dp.data_ptr(), dz.data_ptr(), p.shape[0], p.shape[1], stream_ptr()),
          "softmax_bwd")
    return dz


def add(a, b, relu=False, out=None):
    L = load()
    y = out if out is not None else torch.empty_like(a)
    check(L.tfde_add(_f32c(a, "a").data_ptr(), _f32c(b, "b").data_ptr(), y.data_ptr(), a.numel(),
                     1 if relu else 0, stream_ptr()), "add")
    return y


def u8_to_f32(x, scale=1.0 / 255.0, out=None):
    L = load()
    y = out if out is not None else torch.empty(x.shape, dtype=torch.float32, device=x.device)
    check(L.tfde_u8_to_f32(x.data_ptr(), y.data_ptr(), x.numel(), scale, stream_ptr()), "u8_to_f32")
    return y


def dropout(x, rate, hparams, stream_id, training=True, out=None):
    """Forward and backward are the same call (the mask is regenerated from the device-side step)."""
    L = load()
    y = out if out is not None else torch.empty_like(x)
    check(L.tfde_dropout(_f32c(x, "x").data_ptr(), y.data_ptr(), x.numel(), rate, hparams.data_ptr(),
                         stream_id, 1 if training else 0, stream_ptr()), "dropout")
    return y


def global_avgpool_fwd(x, out=None):
    L = load()
    bn, h, wd, c = x.shape
    y = out if out is not None else torch.empty((bn, c), dtype=torch.float32, device=x.device)
    check(L.tfde_gap_fwd(_f32c(x, "x").data_ptr(), y.data_ptr(), bn, h * wd, c, stream_ptr()), "gap_fwd")
    return y


def global_avgpool_bwd(dy, x_shape, out=None):
    L = load()
    bn, h, wd, c = x_shape
    dx = out if out is not None else torch.empty(tuple(x_shape), dtype=torch.float32, device=dy.device)
    check(L.tfde_gap_bwd(_f32c(dy, "dy").data_ptr(), dx.data_ptr(), bn, h * wd, c, stream_ptr()), "gap_bwd")
    return dx


# ----------------------------------------------------------------------------------------------
# loss / optimizer
# ----------------------------------------------------------------------------------------------
def xent(inp, labels, scale, from_probs=False, out_grad=None, out4=None):
    """Returns (out4, dgrad): out4 = [loss_sum*scale, correct, rows, 0] on device."""
    L = load()
    _f32c(inp, "logits")
    rows, c = inp.shape
    labels = labels.reshape(-1)
    if labels.dtype not in (torch.int64, torch.int32):
        raise TypeError("labels must be int32/int64")
    if not labels.is_contiguous():
        labels = labels.contiguous()
    dev = inp.device
    wsp = Workspace.current(dev)
    dg = out_grad if out_grad is not None else torch.empty_like(inp)
    o4 = out4 if out4 is not None else torch.empty((4,), dtype=torch.float32, device=dev)
    check(L.tfde_xent(inp.data_ptr(), labels.data_ptr(), 1 if labels.dtype == torch.int64 else 0,
                      dg.data_ptr(), rows, c, scale, 1 if from_probs else 0, wsp.small.data_ptr(),
                      wsp.counters.data_ptr(), o4.data_ptr(), stream_ptr()), "xent")
    return o4, dg


def optimizer_apply(w, g, hparams, opt=OPT_SGD, s1=None, s2=None, w_bf16=None, n_extra=0, metrics=None,
                    last_extra=None):
    """In-place fused update of the flat arena ``w`` from ``g`` (g has n + n_extra floats)."""
    L = load()
    wsp = Workspace.current(w.device)
    check(L.tfde_optimizer_apply(w.data_ptr(), g.data_ptr(), ptr(s1), ptr(s2), ptr(w_bf16), w.numel(),
                                 n_extra, ptr(metrics), ptr(last_extra), opt, hparams.data_ptr(),
                                 wsp.counters.data_ptr() + 4 * (MAX_COUNTERS - 1), stream_ptr()),
          "optimizer_apply")


# ----------------------------------------------------------------------------------------------
# bf16 tensor-core path (csrc/kernels/gemm_tcgen05.cu)
# ----------------------------------------------------------------------------------------------
register_optional("tfde_gemm_bf16_tn", [c_p, c_p, c_p, c_p, c_l, c_p, c_i, c_i, c_i, c_i, c_p])
register_optional("tfde_cast_bf16", [c_p, c_p, c_l, c_p])
register_optional("tfde_cast_transpose_bf16", [c_p, c_p, c_p, c_i, c_i, c_p])
register_optional("tfde_dense_fwd_static_scratch", [c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_p])


def gemm_bf16_tn(a, b, bias=None, relu=False, out_dtype=torch.float32, out=None):
    """C[M,N] = act(A[M,K] @ B[N,K]^T + bias) on tcgen05 tensor cores (TMEM accumulator, TMA-fed).
    ``a``/``b`` are bf16, row-major, K contiguous, K % 8 == 0."""
    L = load()
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
    check(L.tfde_gemm_bf16_tn(a.data_ptr(), b.data_ptr(), f32, b16, c.stride(0), ptr(bias), 1 if relu else 0,
                              M, N, K, stream_ptr()), "gemm_bf16_tn")
    return c


def cast_bf16(x, out=None):
    L = load()
    y = out if out is not None else torch.empty(x.shape, dtype=torch.bfloat16, device=x.device)
    check(L.tfde_cast_bf16(_f32c(x, "x").data_ptr(), y.data_ptr(), x.numel(), stream_ptr()), "cast_bf16")
    return y


def cast_transpose_bf16(x, mask=None, out=None):
    """y[C,R] (bf16) = (x * (mask > 0))^T for fp32 x[R,C]."""
    L = load()
    R, C = x.shape
    y = out if out is not None else torch.empty((C, R), dtype=torch.bfloat16, device=x.device)
    check(L.tfde_cast_transpose_bf16(_f32c(x, "x").data_ptr(), ptr(mask), y.data_ptr(), R, C, stream_ptr()),
          "cast_transpose_bf16")
    return y


def dense_fwd_static_scratch(x, w, bias=None, relu=False, out=None):
    """Dense forward whose split-K scratch is the library's built-in static device array (no allocation
    besides the output unless ``out`` is given). See include/tfde_kernels.h for the limits."""
    L = load()
    _f32c(x, "x"); _f32c(w, "w")
    M, K = x.shape
    N = w.shape[1]
    y = out if out is not None else torch.empty((M, N), dtype=torch.float32, device=x.device)
    check(L.tfde_dense_fwd_static_scratch(x.data_ptr(), w.data_ptr(), ptr(bias), y.data_ptr(), M, K, N,
                                          1 if relu else 0, stream_ptr()), "dense_fwd_static_scratch")
    return y


# ----------------------------------------------------------------------------------------------
# small-batch latency kernels (csrc/kernels/small_batch.cu)
# ----------------------------------------------------------------------------------------------
_SKINNY = os.environ.get("TFDE_SKINNY", "1") == "1"
register_optional("tfde_dense_skinny_supported", [c_i, c_i, c_i])
register_optional("tfde_dense_skinny_ws_floats", [c_i, c_i, c_i], c_z)
register_optional("tfde_dense_skinny_fwd", [c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_p, c_p, c_p])
register_optional("tfde_head_supported", [c_i, c_i])
register_optional("tfde_head_ws_floats", [c_i, c_i, c_i], c_z)
register_optional("tfde_head_fwd_bwd", [c_p, c_p, c_p, c_p, c_i, c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_i,
                                        c_f, c_p, c_p, c_p])


def head_supported(K: int, N: int) -> bool:
    return bool(load().tfde_head_supported(K, N))


def head_fwd_bwd(h, w, bias, labels, scale, from_probs=False, relu_mask=False, need_dh=True, out=None,
                 out_dh=None, out_dw=None, out_db=None, out4=None):
    """Fused Dense(N<=32) + sparse-CE (+accuracy) + backward of that Dense. Returns (out4, out, dh, dw, db)."""
    L = load()
    _f32c(h, "h"); _f32c(w, "w")
    M, K = h.shape
    N = w.shape[1]
    dev = h.device
    labels = labels.reshape(-1)
    if labels.dtype not in (torch.int64, torch.int32):
        raise TypeError("labels must be int32/int64")
    wsp = Workspace.current(dev)
    if L.tfde_head_ws_floats(M, K, N) > wsp.ws.numel():
        raise RuntimeError("head workspace too small")
    dh = (out_dh if out_dh is not None else torch.empty((M, K), dtype=torch.float32, device=dev)) if need_dh else None
    dw = out_dw if out_dw is not None else torch.empty((K, N), dtype=torch.float32, device=dev)
    db = None
    if bias is not None:
        db = out_db if out_db is not None else torch.empty((N,), dtype=torch.float32, device=dev)
    o4 = out4 if out4 is not None else torch.empty((4,), dtype=torch.float32, device=dev)
    check(L.tfde_head_fwd_bwd(h.data_ptr(), w.data_ptr(), ptr(bias), labels.data_ptr(),
                              1 if labels.dtype == torch.int64 else 0, ptr(out), ptr(dh), dw.data_ptr(), ptr(db),
                              o4.data_ptr(), M, K, N, 1 if relu_mask else 0, 1 if from_probs else 0, scale,
                              wsp.ws.data_ptr(), wsp.counters.data_ptr(), stream_ptr()), "head_fwd_bwd")
    return o4, out, dh, dw, db
