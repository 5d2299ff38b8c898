"""sb_cumsum_i32_static -- the prefix sum for callers that may not allocate (include/splat_b200.h; no counterpart in
the reference) -- against the existing entry point sb_cumsum_i32, called
through the C-ABI.  Integer work: the bar is bit-exact.  The file is named so that pytest collects it last."""
import ctypes
import os
import re

import pytest
import torch

from splatformer_b200 import _abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MAX = _abi.CUMSUM_STATIC_MAX_COUNT


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def test_static_entry_point_rejects_sizes_beyond_its_scratch_without_a_device(sb_lib):
    """Argument checks happen before any CUDA call, so they run on the CPU box too."""
    assert sb_lib.sb_cumsum_i32_static(MAX + 1, None, None, None) == _abi.ERR_STATIC_SCRATCH_EXCEEDED
    assert sb_lib.sb_cumsum_i32_static(2 ** 31 - 1, None, None, None) == _abi.ERR_STATIC_SCRATCH_EXCEEDED
    assert sb_lib.sb_cumsum_i32_static(-1, None, None, None) == _abi.ERR_BAD_ARG
    assert sb_lib.sb_cumsum_i32_static(0, None, None, None) == 0
    assert sb_lib.sb_cumsum_i32_static(5, None, None, None) == _abi.ERR_BAD_ARG          # NULL in / out
    msg = sb_lib.sb_error_string(_abi.ERR_STATIC_SCRATCH_EXCEEDED).decode()
    assert "static scratch" in msg and str(MAX) in msg
    with pytest.raises(RuntimeError, match="static scratch"):
        _abi.check(sb_lib.sb_cumsum_i32_static(MAX + 1, None, None, None), "sb_cumsum_i32_static")
    header = open(os.path.join(ROOT, "include", "splat_b200.h")).read()
    m = re.search(r"#define SB_CUMSUM_STATIC_MAX_COUNT \(1 << (\d+)\)", header)
    assert m and (1 << int(m.group(1))) == MAX


def test_static_entry_point_allocates_nothing_in_its_source():
    src = open(os.path.join(ROOT, "splatformer_b200", "csrc", "radix_sort.cu")).read()
    body = src[src.index('extern "C" int sb_cumsum_i32_static'):]
    helper = src[src.index("static int static_scan_scratch"):src.index('extern "C" int sb_cumsum_i32_static')]
    for text in (body, helper):
        code = re.sub(r"//.*", "", text)
        for banned in ("cudaMalloc", "cudaFree", "cudaHostAlloc", "cudaMallocAsync", "new ", "malloc("):
            assert banned not in code, banned
    assert "cudaGetSymbolAddress" in helper
    assert "static __device__ __align__(256)" in src


def _existing(lib, x, stream):
    out = torch.empty_like(x)
    nbytes = lib.sb_cumsum_temp_bytes(x.numel())
    temp = torch.empty(nbytes, dtype=torch.uint8, device=x.device)
    _abi.check(lib.sb_cumsum_i32(x.numel(), _ptr(x), _ptr(out), _ptr(temp), nbytes, stream), "sb_cumsum_i32")
    return out


def _static(lib, x, out, stream):
    _abi.check(lib.sb_cumsum_i32_static(x.numel(), _ptr(x), _ptr(out), stream), "sb_cumsum_i32_static")
    return out


@pytest.mark.gpu
def test_static_cumsum_matches_the_existing_entry_point_on_calls_in_a_row(cuda, sb_lib):
    """Two (and more) calls in a row share the static scratch, which keeps the previous call's look-back state:
    large after small, small after large, the same size twice, ragged and unaligned inputs, the maximum size."""
    with torch.cuda.device(cuda):
        stream = _abi.stream()
        g = torch.Generator(device="cpu").manual_seed(1234)
        sizes = [100_003, 100_003, 1, 4095, 4096, 4097, 3_200_000, 33, 8 * 4096 + 3, MAX, 5000, MAX]
        for call, count in enumerate(sizes):
            if count >= 1 << 22:
                x = torch.randint(0, 100, (count,), dtype=torch.int32, device=cuda)         # sum < 2^31
            else:
                x = torch.randint(0, 50, (count,), generator=g, dtype=torch.int32).to(cuda)
            out = torch.full((count,), -7, dtype=torch.int32, device=cuda)
            _static(sb_lib, x, out, stream)
            ref = _existing(sb_lib, x, stream)
            assert torch.equal(out, ref), f"call {call}, count {count}: differs from sb_cumsum_i32"
            assert torch.equal(out, torch.cumsum(x, 0, dtype=torch.int32)), f"call {call}, count {count}"
        # an input / output that is not 16-byte aligned (the scan's scalar path), twice in a row
        for _ in range(2):
            buf = torch.randint(0, 50, (70_001,), generator=g, dtype=torch.int32).to(cuda)
            dst = torch.zeros(70_001, dtype=torch.int32, device=cuda)
            _static(sb_lib, buf[1:], dst[1:], stream)
            assert torch.equal(dst[1:], _existing(sb_lib, buf[1:].contiguous(), stream))
        torch.cuda.synchronize()


@pytest.mark.gpu
def test_static_cumsum_leaves_the_allocators_alone_and_runs_on_another_stream(cuda, sb_lib):
    with torch.cuda.device(cuda):
        x = torch.randint(0, 50, (1_000_000,), dtype=torch.int32, device=cuda)
        out = torch.empty_like(x)
        _static(sb_lib, x, out, _abi.stream())                  # set-up call: symbol address, module load
        ref = _existing(sb_lib, x, _abi.stream())
        torch.cuda.synchronize()
        side = torch.cuda.Stream()
        torch_before = torch.cuda.memory_stats(cuda)["allocation.all.allocated"]
        free_before, _ = torch.cuda.mem_get_info(cuda)
        with torch.cuda.stream(side):
            for _ in range(2):
                out.fill_(-1)
                _static(sb_lib, x, out, ctypes.c_void_p(side.cuda_stream))
        side.synchronize()
        free_after, _ = torch.cuda.mem_get_info(cuda)
        assert torch.cuda.memory_stats(cuda)["allocation.all.allocated"] == torch_before
        assert free_after >= free_before, (free_before, free_after)
        assert torch.equal(out, ref)
        # beyond the maximum: refused, nothing launched, the output untouched
        out.fill_(-1)
        rc = sb_lib.sb_cumsum_i32_static(MAX + 1, _ptr(x), _ptr(out), _abi.stream())
        assert rc == _abi.ERR_STATIC_SCRATCH_EXCEEDED
        torch.cuda.synchronize()
        assert bool((out == -1).all())
