"""The C-ABI library loads on a machine without a GPU and exports every symbol that
include/splat_b200.h declares, with the ctypes signatures of splatformer_b200/_abi.py in step."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "splat_b200.h")


def _declared():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    out = {}
    for m in re.finditer(r"SB_API\s+([\w\s\*]+?)\s*\**\s*(sb_\w+)\s*\(([^;]*?)\)\s*;", text, flags=re.S):
        args = m.group(3).strip()
        nargs = 0 if args in ("", "void") else len([a for a in args.split(",") if a.strip()])
        out[m.group(2)] = nargs
    return out


def test_header_declares_the_expected_entry_points():
    d = _declared()
    for name in ("sb_sh_fwd", "sb_sh_bwd", "sb_project_fwd", "sb_project_bwd", "sb_cumsum_i32",
                 "sb_map_gaussian_to_intersects", "sb_sort_pairs", "sb_tile_bin_edges", "sb_rasterize_fwd",
                 "sb_rasterize_bwd", "sb_views_project_fwd", "sb_views_project_bwd", "sb_sort_tiles"):
        assert name in d, name


def test_library_exports_every_declared_symbol(sb_lib):
    from splatformer_b200 import _abi
    d = _declared()
    raw = ctypes.CDLL(_abi.lib_path())
    for name, nargs in d.items():
        assert hasattr(raw, name), f"{name} declared in the header but not exported"
        assert name in _abi.SIGNATURES, f"{name} has no ctypes signature"
        assert len(_abi.SIGNATURES[name][1]) == nargs, f"{name}: header has {nargs} args, binding " \
                                                       f"{len(_abi.SIGNATURES[name][1])}"
    for name in _abi.SIGNATURES:
        assert name in d, f"{name} bound but not declared in include/splat_b200.h"
    assert sb_lib.sb_abi_version() == _abi.ABI_VERSION
    assert b"invalid argument" in sb_lib.sb_error_string(-1)


def test_no_torch_types_or_cpu_fallback_in_the_boundary():
    text = open(HEADER).read()
    assert "torch" not in text.lower().replace("pytorch", "") or "no torch types" in text
    assert "at::Tensor" not in text
    # host-side argument validation works without a device (no compute is launched)
    from splatformer_b200 import _abi
    lib = _abi.load()
    assert lib.sb_sort_pairs(-1, 40, None, None, None, None, None, 0, None) == -1
    assert lib.sb_rasterize_fwd(1, 0, None, None, None, None, 16, 16, 17, None, None, None, None, None, 0, None, 0,
                                None, None) == -1           # block_width 17
    assert lib.sb_sort_tiles(1, 64, 100, None, None, None, None, None, None, 0, 0, None, None, None, None, 0, None) == -1    # capacity not a power of two
    assert lib.sb_sh_fwd(1, 4, 2, 3, None, None, None, None) == -1      # degrees_to_use > degree


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "splatformer_b200")
    for base, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(base, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", src, flags=re.M), f"{f} imports the oracle"
    for f in os.listdir(os.path.join(ROOT, "gsplat")):
        if f.endswith(".py"):
            assert "oracle" not in open(os.path.join(ROOT, "gsplat", f)).read()


def test_cpu_tensors_are_rejected():
    import torch
    import gsplat
    with pytest.raises(RuntimeError):
        gsplat.spherical_harmonics(0, torch.zeros(4, 3), torch.zeros(4, 1, 3))
    from splatformer_b200 import render
    with pytest.raises(RuntimeError):
        render.render_views(torch.zeros(4, 3), torch.ones(4, 3), torch.ones(4, 4), torch.ones(4, 1),
                            torch.zeros(4, 3), -1, torch.eye(4)[None, :3], None, torch.ones(1, 4), 16, 16)
