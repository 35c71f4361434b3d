"""Bring-up aid for the implicit-GEMM convolutions: per case, the error of forward / data gradient / filter gradient
against an fp32 PyTorch reference, and where on the image the wrong elements sit when there are any."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch

from test_conv_igemm_gpu import CASES, _ref, _setup   # noqa: E402


def where(tag, got, ref, tol):
    err = (got - ref).abs()
    scale = ref.abs().max().item()
    bad = err > tol * scale
    msg = f"  {tag}: max err {err.max().item():.4g} / scale {scale:.4g}; wrong {int(bad.sum())} of {bad.numel()}"
    if bad.any():
        idx = bad.nonzero()[0].tolist()
        msg += f"; first wrong index {idx}: got {got[tuple(idx)].item():.4g} want {ref[tuple(idx)].item():.4g}"
        if got.dim() == 4 and got.shape[1] <= 32:
            grid = bad[0].any(dim=-1).int()
            msg += "\n    image 0 wrong-pixel map:\n" + "\n".join("      " + "".join(".#"[v] for v in row) for row in grid.tolist())
        if torch.isnan(got).any():
            msg += f"\n    NaNs: {int(torch.isnan(got).sum())}"
    print(msg, flush=True)
    return not bad.any()


def main():
    from tfde_b200.ops import conv_igemm
    ok = True
    for case in CASES:
        print("case", case, flush=True)
        x, w, dy, stride, pads, ho, wo = _setup(case)
        xr, wr = x.float().requires_grad_(True), w.float().requires_grad_(True)
        ref = _ref(xr, wr, stride, pads)
        ref.backward(dy.float())
        y = torch.empty(x.shape[0], ho, wo, w.shape[3], dtype=torch.bfloat16, device="cuda")
        conv_igemm.fwd(x.cuda(), w.cuda(), None, stride, pads, y)
        torch.cuda.synchronize()
        ok &= where("fwd  ", y.float().cpu(), ref.detach(), 1e-2)
        dw = torch.zeros(tuple(w.shape), dtype=torch.float32, device="cuda")
        conv_igemm.wgrad(x.cuda(), dy.cuda(), tuple(w.shape), stride, pads, dw)
        torch.cuda.synchronize()
        ok &= where("wgrad", dw.cpu(), wr.grad, 2e-4)
        if stride == (1, 1):
            dx = torch.empty(tuple(x.shape), dtype=torch.bfloat16, device="cuda")
            conv_igemm.dgrad(dy.cuda(), w.cuda(), pads, dx)
            torch.cuda.synchronize()
            ok &= where("dgrad", dx.float().cpu(), xr.grad, 1e-2)
    print("ALL OK" if ok else "MISMATCH", flush=True)
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
