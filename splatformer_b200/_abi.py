"""ctypes binding of libsplat_b200.so (the C ABI declared in include/splat_b200.h).

There is no fallback of any kind: if the library is missing or a CUDA tensor is not what the ABI
expects, the call raises.  The library is built in-tree by `python -m splatformer_b200.build`
(`__graft_entry__.build()` does that).
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_float, c_int, c_size_t, c_void_p
from typing import Optional

import torch

_LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib", "libsplat_b200.so")
_lib: Optional[ctypes.CDLL] = None

P, I, F, S = c_void_p, c_int, c_float, c_size_t
ABI_VERSION = 3

# input-convention flags of sb_views_project_fwd / _bwd (include/splat_b200.h SB_FLAG_*)
FLAG_RAW_SCALES, FLAG_RAW_OPACITIES, FLAG_COLOR_SIGMOID, FLAG_C2W_OPENGL = 1, 2, 4, 8
FLAG_INTRINS_SHARED = 16            # intrins is (1,4), shared by all views
FLAG_RAW_QUATS = 32                 # quats are raw: wrapper's q / |q| and (0,0,0,1) substitution in-kernel
ERR_BAD_ARG, ERR_TEMP_TOO_SMALL, ERR_STATIC_SCRATCH_EXCEEDED = -1, -2, -3
CUMSUM_STATIC_MAX_COUNT = 1 << 24   # sb_cumsum_i32_static: elements its file-scope device scratch covers

# name -> (restype, argtypes); mirrors include/splat_b200.h one to one
SIGNATURES = {
    "sb_abi_version": (I, []),
    "sb_error_string": (c_char_p, [I]),
    "sb_sh_fwd": (I, [I, I, I, I, P, P, P, P]),
    "sb_sh_bwd": (I, [I, I, I, I, P, P, P, P]),
    "sb_project_fwd": (I, [I, I, P, P, F, P, P, I, P, F, F, F, F, I, I, I, F, P, P, P, P, P, P, P, P]),
    "sb_project_bwd": (I, [I, I, P, P, F, P, P, I, P, F, F, F, F, I, I, I, F, P, P, P, P, P, P, P, P, P, P, P]),
    "sb_cumsum_temp_bytes": (S, [I]),
    "sb_cumsum_i32": (I, [I, P, P, P, S, P]),
    "sb_cumsum_i32_static": (I, [I, P, P, P]),
    "sb_map_gaussian_to_intersects": (I, [I, I, P, I, P, P, P, I, I, I, P, P, P]),
    "sb_depth_order_keys": (I, [I, I, P, P, P, P, P]),
    "sb_gather_i32": (I, [I, P, P, P, P]),
    "sb_map_gaussian_to_intersects_ordered": (I, [I, I, P, I, P, P, P, P, I, I, I, P, P, P]),
    "sb_sort_pairs_range": (I, [I, I, I, P, P, P, P, P, S, P]),
    "sb_map_gaussian_to_tiles_ordered": (I, [I, I, P, I, P, P, P, P, I, I, I, P, P, P]),
    "sb_sort_tiles_global_temp_bytes": (S, [I, I]),
    "sb_sort_tiles_global": (I, [I, I, P, P, P, P, P, P, S, P]),
    "sb_sort32_temp_bytes": (S, [I]),
    "sb_sort_pairs32": (I, [I, I, P, P, P, P, P, S, P]),
    "sb_tile_bin_edges32": (I, [I, P, I, P, P]),
    "sb_sort_temp_bytes": (S, [I]),
    "sb_sort_pairs": (I, [I, I, P, P, P, P, P, S, P]),
    "sb_tile_bin_edges": (I, [I, P, I, P, P]),
    "sb_pack_splats": (I, [I, P, P, P, P, P, P]),
    "sb_rasterize_fwd": (I, [I, I, P, P, P, P, I, I, I, P, P, P, P, P, I, P, I, P, P]),
    "sb_rasterize_bwd": (I, [I, I, P, P, P, P, I, I, I, P, P, P, P, P, I, P, I, P, P, P]),
    "sb_unpack_splat_grads": (I, [I, P, P, P, P, P, P, P, P]),
    "sb_cov2d_bounds": (I, [I, P, P, P, P]),
    "sb_views_project_fwd": (I, [I, I, P, P, F, P, P, P, P, I, I, P, I, P, F, F, F, F, P, I, I, I, F, P, P, P, P,
                                 P, P, I, P]),
    "sb_tile_sort_capacity": (I, []),
    "sb_bin_to_tile_segments": (I, [I, I, P, I, P, P, I, I, I, P, P, I, P]),
    "sb_sort_tiles": (I, [I, I, I, P, P, P, P, P, P, I, I, P, P, P, P, I, P]),
    "sb_rasterize_rect_fwd": (I, [I, I, P, P, P, P, I, I, P, P, P, P, P, I, P, I, P, P, P]),
    "sb_rasterize_rect_bwd": (I, [I, I, P, P, P, P, I, I, P, P, P, P, P, I, P, I, P, P, P]),
    "sb_views_project_bwd": (I, [I, I, P, P, F, P, I, I, P, I, P, F, F, F, F, P, I, I, I, F, P, P, P, P, P, P, P,
                                 P, P, P]),
}

# number of ABI calls that enqueued device work since import (bench.py reports it as gpu_launches)
LAUNCHES = 0


def lib_path() -> str:
    return _LIB_PATH


def load() -> ctypes.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB_PATH):
        raise ImportError(
            f"{_LIB_PATH} is missing: the CUDA extension has not been built. Run "
            "`python -m splatformer_b200.build` (needs nvcc). There is no CPU fallback.")
    lib = ctypes.CDLL(_LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export it
        fn.restype = res
        fn.argtypes = args
    if lib.sb_abi_version() != ABI_VERSION:
        raise ImportError(f"libsplat_b200.so ABI version {lib.sb_abi_version()} != {ABI_VERSION}; rebuild")
    _lib = lib
    return lib


def check(code: int, what: str) -> None:
    global LAUNCHES
    LAUNCHES += 1
    if code != 0:
        msg = load().sb_error_string(code).decode()
        raise RuntimeError(f"{what} failed: {msg} (code {code})")


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def stream() -> c_void_p:
    """The current stream of the current device as the ABI's sb_stream_t.  The raw lookup is ~1 us; going
    through a torch.cuda.Stream object costs ~20 us per call, which matters to the wrapper path's
    ~12 launches per view."""
    if _raw_stream is not None:
        return c_void_p(_raw_stream(torch.cuda.current_device()))
    return c_void_p(torch.cuda.current_stream().cuda_stream)


def ptr(t: Optional[torch.Tensor], dtype: torch.dtype, name: str) -> c_void_p:
    """Device pointer of a contiguous CUDA tensor of the expected dtype (None -> NULL)."""
    if t is None:
        return c_void_p(None)
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (the renderer has no CPU path)")
    if t.dtype != dtype:
        raise TypeError(f"{name} must have dtype {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise RuntimeError(f"{name} must be contiguous")
    return c_void_p(t.data_ptr())


def aligned16(t: torch.Tensor) -> torch.Tensor:
    """float4 paths need 16-byte aligned bases; a sliced view can break that."""
    return t if t.data_ptr() % 16 == 0 else t.clone()
