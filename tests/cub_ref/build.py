"""Builds tests/cub_ref/libcub_ref.so (test-only CUB reference, see cub_ref.cu)."""
import ctypes
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "libcub_ref.so")
SRC = os.path.join(HERE, "cub_ref.cu")


def build(force: bool = False) -> str:
    if not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= os.path.getmtime(SRC):
        return LIB
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    cmd = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-shared", "-Xcompiler", "-fPIC",
           "-Xcompiler", "-fvisibility=hidden", SRC, "-o", LIB + ".tmp", "-lcudart"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"nvcc failed for cub_ref.cu:\n{res.stdout}\n{res.stderr}")
    os.replace(LIB + ".tmp", LIB)
    return LIB


def load() -> ctypes.CDLL:
    lib = ctypes.CDLL(build())
    P, I, S = ctypes.c_void_p, ctypes.c_int, ctypes.c_size_t
    for name in ("cubref_sort64_temp", "cubref_sort32_temp", "cubref_scan_temp"):
        getattr(lib, name).restype, getattr(lib, name).argtypes = S, [I]
    for name in ("cubref_sort64", "cubref_sort32"):
        getattr(lib, name).restype, getattr(lib, name).argtypes = I, [I, I, I, P, P, P, P, P, S, P]
    lib.cubref_scan.restype, lib.cubref_scan.argtypes = I, [I, P, P, P, S, P]
    return lib


if __name__ == "__main__":
    print(build(force=True))
