"""Build libconceptattention_b200.so in-tree with nvcc for sm_100a.

    python -m conceptattention_b200.build [--force]

nvcc cross-compiles without a GPU; the .so is git-ignored but travels to the GPU
box with the snapshot.  Nothing here touches torch: the library's ABI is plain C
(include/conceptattention_b200.h).
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

PKG = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG, "csrc")
OBJ = os.path.join(PKG, "build")
LIB = os.path.join(PKG, "libconceptattention_b200.so")
SOURCES = ["ca_api.cu", "k1_prep.cu", "k2_attention.cu", "k3_heatmap.cu", "k3_heatmap_tc.cu", "k5_masks.cu",
           "k6_metrics.cu", "k7_seqpar.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
         "-Xcompiler", "-fPIC", "-Xptxas", "-v"]
# Only K2 is built with --use_fast_math (its exp2 / reciprocal are explicit approximations chosen in
# the source, and flush-to-zero is wanted there); the HBM-bound kernels (K1, K3-K7) keep IEEE
# division, sqrt and denormals so that their results do not depend on a compiler flag.
EXTRA_FLAGS = {"k2_attention.cu": ["--use_fast_math"]}


def _deps():
    hdrs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    hdrs.append(os.path.join(os.path.dirname(PKG), "include", "conceptattention_b200.h"))
    return hdrs


def _stale(target, srcs):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in srcs)


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJ, exist_ok=True)
    deps = _deps()

    def compile_one(src):
        s, o = os.path.join(CSRC, src), os.path.join(OBJ, src.replace(".cu", ".o"))
        if force or _stale(o, [s] + deps):
            r = subprocess.run([NVCC, *FLAGS, *EXTRA_FLAGS.get(src, []), "-c", s, "-o", o], capture_output=True, text=True)
            if r.returncode != 0:
                raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
            with open(o + ".ptxas.log", "w") as f:
                f.write(r.stderr)
            if verbose:
                print(r.stderr)
        return o

    with ThreadPoolExecutor(max_workers=len(SOURCES)) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    if force or _stale(LIB, objs):
        r = subprocess.run([NVCC, "-shared", "-o", LIB, *objs, "-lcudart"], capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
