"""Host-side logic of the implicit-GEMM convolution path (no GPU): which layers it takes, what that means for the
im2col cache of the layer, and that the CPU reference path of act16 is unaffected by the switch."""
import torch

from tfde_b200.ops import act16, conv_igemm


def test_supported_geometries():
    ok = conv_igemm.supported
    assert ok((64, 56, 56, 64), (3, 3, 64, 64), 56, 56)
    assert ok((64, 7, 7, 512), (3, 3, 512, 512), 7, 7)
    assert ok((2, 16, 16, 64), (3, 3, 64, 128), 8, 8)            # strided: forward and filter gradient
    assert not ok((64, 56, 56, 64), (1, 1, 64, 256), 56, 56)     # 1x1: the activation is the operand already
    assert not ok((64, 224, 224, 3), (7, 7, 3, 64), 112, 112)    # 6-byte pixels: below the 16-byte TMA granule
    assert not ok((128, 14, 14, 12), (6, 6, 12, 24), 7, 7)       # Model B's layers: direct kernels instead
    assert not ok((8, 8, 8, 64), (3, 3, 64, 24), 8, 8)           # output channels not a multiple of 64
    assert conv_igemm.dgrad_eligible((1, 1)) and not conv_igemm.dgrad_eligible((2, 2))


def test_no_im2col_cache_when_the_operand_is_fetched_by_tma(monkeypatch):
    x_shape, w_shape = (64, 56, 56, 64), (3, 3, 64, 64)
    monkeypatch.setattr(conv_igemm, "ENABLED", True)
    assert act16.a_operand_shape(x_shape, w_shape, (1, 1), 56, 56, True) is None
    monkeypatch.setattr(conv_igemm, "ENABLED", False)
    assert act16.a_operand_shape(x_shape, w_shape, (1, 1), 56, 56, True) == (64 * 56 * 56, 576)
    monkeypatch.setattr(conv_igemm, "ENABLED", True)
    # the fp32 network input (stem) is gathered and cast in one kernel: never the implicit GEMM
    assert act16.a_operand_shape((8, 32, 32, 64), w_shape, (1, 1), 32, 32, False) == (8 * 32 * 32, 576)


def test_cpu_reference_path_is_independent_of_the_switch(monkeypatch):
    g = torch.Generator().manual_seed(0)
    x = torch.randn(2, 6, 6, 64, generator=g).to(torch.bfloat16)
    w = torch.randn(3, 3, 64, 64, generator=g) * 0.05
    outs = []
    for on in (True, False):
        monkeypatch.setattr(conv_igemm, "ENABLED", on)
        y = act16.conv2d_fwd(x, w, None, (1, 1), (1, 1, 1, 1))
        dx, dw, _ = act16.conv2d_bwd(x, w, y, (1, 1), (1, 1, 1, 1), has_bias=False)
        outs.append((y, dx, dw))
    for a, b in zip(*outs):
        assert torch.equal(a, b)
