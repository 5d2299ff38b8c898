"""Builds libsplat_b200.so (the C-ABI extension) in-tree with nvcc for sm_100a.

    python -m splatformer_b200.build [--force] [--verbose]

Every translation unit is compiled separately so that the projection FORWARD can be built with
FMA contraction off (`--fmad=false`, the arithmetic contract in csrc/sb_math.cuh: its outputs decide
the tiling and are bit-identical to the oracle) while everything else -- the rasterisers and the
projection backward, whose outputs are gradients -- keeps FMA.  Objects go to splatformer_b200/lib/obj/, the library to
splatformer_b200/lib/libsplat_b200.so (git-ignored; it travels to the GPU box with gpurun).
"""
from __future__ import annotations

import argparse
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

PKG = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG, "csrc")
LIBDIR = os.path.join(PKG, "lib")
OBJDIR = os.path.join(LIBDIR, "obj")
LIB = os.path.join(LIBDIR, "libsplat_b200.so")
INCLUDE = os.path.join(os.path.dirname(PKG), "include")

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
          "--expt-relaxed-constexpr", "-I", INCLUDE]
# translation unit -> extra flags
UNITS = {
    "abi.cu": [],
    "sh.cu": [],
    "project_fwd.cu": ["--fmad=false"],
    "project_bwd.cu": [],
    "tiles.cu": [],
    "radix_sort.cu": [],
    "tile_sort.cu": [],
    "raster_fwd.cu": [],
    "raster_bwd.cu": [],
    "raster_rect.cu": [],
}


def _nvcc() -> str:
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found: the CUDA extension cannot be built")
    return exe


def _deps_mtime() -> float:
    m = 0.0
    for root in (CSRC, INCLUDE):
        for f in os.listdir(root):
            if f.endswith((".cu", ".cuh", ".h")):
                m = max(m, os.path.getmtime(os.path.join(root, f)))
    return max(m, os.path.getmtime(os.path.abspath(__file__)))


def needs_build() -> bool:
    return not os.path.exists(LIB) or os.path.getmtime(LIB) < _deps_mtime()


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return LIB
    os.makedirs(OBJDIR, exist_ok=True)
    nvcc = _nvcc()
    extra = ["-Xptxas", "-v"] if verbose else []
    extra += os.environ.get("SB_NVCC_DEFS", "").split()      # tuning experiments: -DNAME=value ...

    def compile_one(item):
        src, flags = item
        obj = os.path.join(OBJDIR, src.replace(".cu", ".o"))
        cmd = [nvcc, *ARCH, *COMMON, *flags, *extra, "-c", os.path.join(CSRC, src), "-o", obj]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{res.stdout}\n{res.stderr}")
        if verbose:
            sys.stderr.write(f"== {src}\n{res.stderr}\n")
        return obj

    with ThreadPoolExecutor(max_workers=min(8, len(UNITS))) as pool:
        objs = list(pool.map(compile_one, UNITS.items()))
    cmd = [nvcc, *ARCH, "-shared", "-o", LIB + ".tmp", *objs, "-lcudart"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"link failed:\n{res.stdout}\n{res.stderr}")
    os.replace(LIB + ".tmp", LIB)
    return LIB


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    a = ap.parse_args()
    print(build(force=a.force, verbose=a.verbose))
