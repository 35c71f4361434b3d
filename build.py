#!/usr/bin/env python
"""Builds the native libraries of tfde_b200 IN-TREE.

* ``libtfde_kernels.so`` -- every sm_100a CUDA kernel under ``csrc/kernels`` (plain C ABI, loaded
  with ctypes; ``cuobjdump -sass`` on it is one command).
* ``libtfde_runtime.so`` -- the host-side C++ runtime under ``csrc/runtime`` (input pipeline,
  no CUDA dependency, so it also builds and runs on the CPU-only dev box).

Both land in ``tensorflow-distributed-example_b200/_lib/`` (git-ignored, but shipped to the GPU box
by gpurun). nvcc cross-compiles for sm_100a without a GPU.

Usage: ``python build.py [--force] [--verbose] [--ptxas-info]``
"""
from __future__ import annotations

import argparse
import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "tensorflow-distributed-example_b200")
LIB_DIR = os.path.join(PKG, "_lib")
OBJ_DIR = os.path.join(ROOT, "build", "obj")
KERNEL_DIR = os.path.join(ROOT, "csrc", "kernels")
RUNTIME_DIR = os.path.join(ROOT, "csrc", "runtime")

NVCC = os.environ.get("NVCC", shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc")
CXX = os.environ.get("CXX", "g++")
ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_FLAGS = ["-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC", "-Xcompiler",
              "-fvisibility=hidden", "--expt-relaxed-constexpr", "-I", os.path.join(ROOT, "include"),
              "-I", KERNEL_DIR]
CXX_FLAGS = ["-O2", "-std=c++17", "-fPIC", "-fvisibility=hidden", "-pthread", "-Wall",
             "-I", os.path.join(ROOT, "include")]


def _digest(paths, extra=""):
    h = hashlib.sha256(extra.encode())
    for p in sorted(paths):
        with open(p, "rb") as f:
            h.update(p.encode())
            h.update(f.read())
    return h.hexdigest()


def _run(cmd, verbose):
    if verbose:
        print(" ".join(cmd), flush=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("build step failed: " + " ".join(cmd))
    if verbose and (r.stdout or r.stderr):
        print(r.stdout + r.stderr, flush=True)
    return r.stdout + r.stderr


def _sources(d, exts):
    if not os.path.isdir(d):
        return []
    return sorted(os.path.join(d, f) for f in os.listdir(d) if f.endswith(exts))


def build_kernels(force=False, verbose=False, ptxas_info=False):
    srcs = _sources(KERNEL_DIR, (".cu",))
    hdrs = _sources(KERNEL_DIR, (".cuh", ".h")) + _sources(os.path.join(ROOT, "include"), (".h",))
    out = os.path.join(LIB_DIR, "libtfde_kernels.so")
    stamp = out + ".stamp"
    digest = _digest(srcs + hdrs, " ".join(ARCH_FLAGS + NVCC_FLAGS))
    if (not force and os.path.exists(out) and os.path.exists(stamp)
            and open(stamp).read().strip() == digest):
        return out
    os.makedirs(LIB_DIR, exist_ok=True)
    os.makedirs(OBJ_DIR, exist_ok=True)
    objs = [os.path.join(OBJ_DIR, os.path.basename(s)[:-3] + ".o") for s in srcs]
    extra = ["-Xptxas", "-v"] if ptxas_info else []

    def compile_one(pair):
        s, o = pair
        return _run([NVCC, *ARCH_FLAGS, *NVCC_FLAGS, *extra, "-c", s, "-o", o], verbose or ptxas_info)

    with ThreadPoolExecutor(max_workers=min(8, max(1, len(srcs)))) as ex:
        logs = list(ex.map(compile_one, zip(srcs, objs)))
    if ptxas_info:
        os.makedirs(os.path.join(ROOT, "profiles"), exist_ok=True)
        with open(os.path.join(ROOT, "profiles", "ptxas_info.txt"), "w") as f:
            f.write("\n".join(logs))
    _run([NVCC, *ARCH_FLAGS, "-shared", "-o", out, *objs, "-lcudart_static", "-lcuda", "-ldl",
          "-lrt", "-lpthread"], verbose)
    with open(stamp, "w") as f:
        f.write(digest)
    return out


def build_runtime(force=False, verbose=False):
    srcs = _sources(RUNTIME_DIR, (".cc", ".cpp"))
    if not srcs:
        return None
    hdrs = _sources(RUNTIME_DIR, (".h",)) + _sources(os.path.join(ROOT, "include"), (".h",))
    out = os.path.join(LIB_DIR, "libtfde_runtime.so")
    stamp = out + ".stamp"
    digest = _digest(srcs + hdrs, " ".join(CXX_FLAGS))
    if (not force and os.path.exists(out) and os.path.exists(stamp)
            and open(stamp).read().strip() == digest):
        return out
    os.makedirs(LIB_DIR, exist_ok=True)
    _run([CXX, *CXX_FLAGS, "-shared", "-o", out, *srcs], verbose)
    with open(stamp, "w") as f:
        f.write(digest)
    return out


def build_all(force=False, verbose=False, ptxas_info=False):
    return build_kernels(force, verbose, ptxas_info), build_runtime(force, verbose)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    ap.add_argument("--ptxas-info", action="store_true")
    a = ap.parse_args()
    print(build_all(a.force or a.ptxas_info, a.verbose, a.ptxas_info))
