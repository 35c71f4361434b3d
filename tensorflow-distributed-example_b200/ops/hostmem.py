"""Zero-copy views of pinned host memory for kernels.

Under unified virtual addressing a ``cudaHostAlloc``'ed (pinned) buffer is directly addressable from the
GPU at the same pointer. ``device_alias`` wraps such a buffer as a CUDA tensor so the hand-written kernels
can read a step's labels from it and write the step's metrics into it without memcpy nodes in the captured
graph (each memcpy node costs ~6 us of the ~70 us step; a 512-byte label read over PCIe costs nothing)."""
from __future__ import annotations

import torch

from . import native

native.register_optional("tfde_copy_small", [native.c_p, native.c_p, native.c_i, native.c_p])


class _HostPtr:
    def __init__(self, t: torch.Tensor):
        self._keep = t
        self.__cuda_array_interface__ = {"shape": (t.numel() * t.element_size(),), "typestr": "|u1",
                                         "data": (t.data_ptr(), False), "version": 2, "strides": None}


def device_alias(pinned: torch.Tensor, device) -> torch.Tensor:
    """CUDA tensor aliasing the pinned host tensor's memory (same dtype/shape)."""
    if not pinned.is_pinned():
        raise ValueError("device_alias needs a pinned host tensor")
    raw = torch.as_tensor(_HostPtr(pinned), device=device)
    return raw.view(pinned.dtype).view(pinned.shape)


def copy_small(dst: torch.Tensor, src: torch.Tensor):
    L = native.load()
    native.check(L.tfde_copy_small(dst.data_ptr(), src.data_ptr(), src.numel(), native.stream_ptr()), "copy_small")
