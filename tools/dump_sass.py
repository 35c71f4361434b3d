"""Writes profiles/sass/<kernel>.sass for the kernels of libtfde_kernels.so and a summary of the SASS
mnemonics that prove the Blackwell-native paths (UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG = TMA load,
UTMALDG.4D.IM2COL = im2col-mode TMA load, UTMASTG = TMA store, UTCBAR = tcgen05.commit, SYNCS = mbarrier, peer/system-scope memory ops in the all-reduce)."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "tensorflow-distributed-example_b200", "_lib", "libtfde_kernels.so")
OUT = os.path.join(ROOT, "profiles", "sass")
KEEP = ("fused_step_kernel", "dense_bwd_skinny_kernel", "ps_push_multi_kernel", "smallconv_dgrad_kernelILi6", "umma_tf32_selftest", "gemm_bf16_tcgen05_kernel", "allreduce_fused_kernel", "dense_skinny_fwd_kernel", "head_fwd_bwd_kernel",
        "conv3x3c1_relu_pool", "allreduce_nvls_kernel", "gemm_bf16_tcgen05_persistent_kernel", "colstats_bf16_kernel",
        "bn_bwd_apply_bf16_kernel", "gemm_fp32_kernel", "smallconv", "bn_reduce_smallc")
N_LISTED = 14     # kernels of KEEP[:N_LISTED] get a full listing; the rest only a row in the summary
MNEMONICS = ("UTCHMMA", "UTCBAR", "LDTM", "UTMALDG", "UTMALDG.4D.IM2COL", "UTMASTG", "UTMAPF", "SYNCS", "LDGSTS", "UCGABAR", "LDGMC", "HMMA", "FFMA", "LDG",
             "STG", "STS", "LDS", "ATOMG", "REDG", "MEMBAR", "CCTL")


def main():
    os.makedirs(OUT, exist_ok=True)
    txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    parts = re.split(r"\n\s*Function : ", txt)
    rows = []
    for part in parts[1:]:
        name, _, body = part.partition("\n")
        if not any(k in name for k in KEEP):
            continue
        dem = subprocess.run(["cu++filt", name.strip()], capture_output=True, text=True).stdout.strip() or name.strip()
        clean = dem.replace("(anonymous namespace)::", "").replace("<unnamed>::", "")
        clean = re.sub(r"^void\s+", "", clean)
        short = re.sub(r"[^A-Za-z0-9_]+", "_", clean.split("(")[0]).strip("_")[:90]
        if "fused_step_kernel" in name:      # two instantiations: exact-fp32 FFMA and TF32 tcgen05
            short = "fused_step_kernel_tf32_tcgen05" if "Lb1" in name else "fused_step_kernel"
        if "smallconv" in name or "bn_reduce_smallc" in name:
            m2 = re.search(r"(smallconv_\w+?_kernel|bn_reduce_smallc_kernel)ILi?b?(\d+)", name)
            if m2:
                short = f"{m2.group(1)}_{m2.group(2)}"
        m3 = re.search(r"(gemm_bf16_tcgen05(?:_persistent)?_kernel|colstats_bf16_kernel|gemm_fp32_kernel)<([^>]*)>", clean)
        if m3:                               # template instantiations: one listing / row each
            short = m3.group(1) + "_" + re.sub(r"[^A-Za-z0-9]+", "_", m3.group(2).replace("(int)", "")).strip("_")
        counts = collections.Counter()
        for line in body.splitlines():
            m = re.search(r"/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
            if m:
                op = m.group(1)
                for mn in MNEMONICS:
                    if op.startswith(mn):
                        counts[mn] += 1
                if ".SYS" in op or ".STRONG.SYS" in op:
                    counts["sys-scope"] += 1
        path = os.path.join(OUT, short + ".sass")
        listed = any(k in name for k in KEEP[:N_LISTED])
        if "persistent_kernel" in short:     # 18 instantiations: list the plain GEMM and the three convolution modes
            listed = short.endswith(("_256_0_0", "_256_2_1", "_256_0_2", "_256_3_3"))
        elif short.startswith("gemm_bf16_tcgen05_kernel"):
            listed = short.endswith("_256_2")
        if listed:
            with open(path, "w") as f:
                f.write("Function : " + part)
        rows.append((short, counts))
    with open(os.path.join(ROOT, "profiles", "sass_summary.md"), "w") as f:
        f.write("# SASS mnemonic counts per kernel (cuobjdump -sass, sm_100a)\n\n")
        f.write("| kernel | " + " | ".join(MNEMONICS) + " | sys-scope |\n|---|" + "---|" * (len(MNEMONICS) + 1) + "\n")
        for short, c in sorted(rows):
            f.write(f"| {short} | " + " | ".join(str(c.get(m, 0)) for m in MNEMONICS) + f" | {c.get('sys-scope', 0)} |\n")
    print(f"wrote {len(rows)} kernels to profiles/sass_summary.md")


if __name__ == "__main__":
    main()
