"""Builds the TEST-ONLY host harness (g++, no CUDA) next to this file and returns a ctypes handle."""
import ctypes
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "harness.cpp")
HDR = os.path.join(HERE, "..", "..", "splatformer_b200", "csrc", "sb_math.cuh")
OUT = os.path.join(HERE, "_hh.so")


def load():
    newest = max(os.path.getmtime(SRC), os.path.getmtime(HDR))
    if not os.path.exists(OUT) or os.path.getmtime(OUT) < newest:
        subprocess.check_call(["g++", "-O2", "-ffp-contract=off", "-fno-fast-math", "-shared", "-fPIC",
                               "-x", "c++", SRC, "-o", OUT])
    return ctypes.CDLL(OUT)
