"""Stand-alone numerics + throughput check of the tcgen05 bf16 GEMM (run under `timeout` on a B200).
Compares against a fp32 PyTorch matmul of the same bf16-rounded operands; prints TFLOP/s per shape and,
for reference, torch.matmul (cuBLAS) bf16 on the same shape."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from tfde_b200.ops import native, tensorcore


def one(M, N, K, bias, relu, out_dtype, time_it=False):
    g = torch.Generator().manual_seed(M * 31 + N * 7 + K)
    a = (torch.randn(M, K, generator=g) * 0.5).to(torch.bfloat16).cuda()
    b = (torch.randn(N, K, generator=g) * 0.5).to(torch.bfloat16).cuda()
    bv = torch.randn(N, generator=g).cuda() if bias else None
    ok_all = True
    for kern in ("tile", "persistent"):
        ck = tensorcore.gemm_bf16_tn(a, b, bv, relu, out_dtype, kernel=kern)
        torch.cuda.synchronize()
        refk = a.float() @ b.float().t()
        if bias:
            refk = refk + bv
        if relu:
            refk = torch.relu(refk)
        errk = (ck.float() - refk).abs().max().item()
        tolk = (2e-3 if out_dtype == torch.float32 else 1.2e-2) * (refk.abs().max().item() + 1e-9)
        if not errk <= tolk:
            print(json.dumps({"kernel": kern, "M": M, "N": N, "K": K, "FAILED_err": errk}), flush=True)
            ok_all = False
    c = tensorcore.gemm_bf16_tn(a, b, bv, relu, out_dtype)
    torch.cuda.synchronize()
    ref = a.float() @ b.float().t()
    if bias:
        ref = ref + bv
    if relu:
        ref = torch.relu(ref)
    err = (c.float() - ref).abs().max().item()
    scale = ref.abs().max().item() + 1e-9
    tol = 2e-3 if out_dtype == torch.float32 else 1.2e-2
    ok = err <= tol * scale and ok_all
    res = {"M": M, "N": N, "K": K, "bias": bias, "relu": relu, "out": str(out_dtype).split(".")[-1],
           "max_abs_err": err, "scale": scale, "ok": bool(ok)}
    if time_it:
        for fn, key in ((lambda: tensorcore.gemm_bf16_tn(a, b, bv, relu, out_dtype, out=c, kernel="tile"), "tile_tflops"),
                        (lambda: tensorcore.gemm_bf16_tn(a, b, bv, relu, out_dtype, out=c, kernel="persistent"), "persistent_tflops"),
                        (lambda: torch.matmul(a, b.t()), "cublas_tflops")):
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(10):
                fn()
            e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 10
            res[key] = 2.0 * M * N * K / (ms * 1e-3) / 1e12
            res[key.replace("tflops", "us")] = ms * 1e3
    print(json.dumps(res), flush=True)
    return ok


def main():
    native.load()
    ok = True
    ok &= one(128, 64, 64, False, False, torch.float32)
    ok &= one(128, 64, 256, False, False, torch.float32)
    ok &= one(128, 128, 512, True, True, torch.float32)
    ok &= one(256, 256, 1024, True, False, torch.float32)
    ok &= one(100, 40, 72, True, True, torch.float32)          # M, N, K tails
    ok &= one(8192, 64, 5408, True, True, torch.float32, True)  # Model A dense at batch 8192
    ok &= one(8192, 5408, 64, False, False, torch.float32, True)  # its dgrad
    ok &= one(5408, 64, 8192, False, False, torch.float32, True)  # its wgrad
    ok &= one(333, 200, 1176, False, False, torch.bfloat16)
    ok &= one(100352, 64, 64, False, True, torch.float32, True)     # ResNet 1x1 conv, 56x56x64, batch 32
    ok &= one(100352, 64, 576, True, False, torch.float32, True)    # ResNet 3x3 conv as im2col GEMM
    ok &= one(6272, 512, 2304, False, False, torch.float32, True)   # stage-4 3x3
    ok &= one(1568, 2048, 512, False, False, torch.float32, True)   # stage-5 1x1 expand
    ok &= one(576, 64, 100352, False, False, torch.float32, True)   # wgrad of a 3x3x64 conv: split-K
    ok &= one(4608, 512, 1568, False, False, torch.float32, True)   # wgrad of the stage-5 3x3 conv
    ok &= one(152, 64, 401408, False, False, torch.float32, True)   # wgrad of conv1 (7x7x3 padded to 152)
    ok &= one(4096, 4096, 4096, False, False, torch.bfloat16, True)
    ok &= one(8192, 8192, 8192, False, False, torch.bfloat16, True)
    print("TCGEN05", "PASSED" if ok else "FAILED", flush=True)
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
