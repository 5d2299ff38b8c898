"""csrc/radix_sort.cu -- the hand-written one-sweep radix sort and single-pass prefix sum -- against
torch.sort(stable=True) / torch.cumsum and, at sizes where those are slow or use the same library underneath,
against the test-only CUB build (tests/cub_ref).  Bit-exact, including the order of equal keys."""
import ctypes

import pytest
import torch

from splatformer_b200 import _abi, utils

pytestmark = pytest.mark.gpu

TILE32, TILE64 = 512 * 16, 512 * 12          # pairs per CTA tile of the two key widths (radix_sort.cu)


def _stable_ref(keys_i64_masked, vals):
    order = torch.sort(keys_i64_masked, stable=True).indices
    return order


@pytest.mark.parametrize("count", [1, 31, 32, 33, 511, TILE32 - 1, TILE32, TILE32 + 1, 3 * TILE32 + 17, 1_000_003])
@pytest.mark.parametrize("bits,dist", [(16, "uniform"), (13, "uniform"), (16, "skewed"), (32, "uniform"), (8, "const")])
def test_sort_pairs32_matches_stable_sort(cuda, sb_lib, count, bits, dist):
    g = torch.Generator(device="cpu").manual_seed(count * 31 + bits)
    if dist == "uniform":
        keys = torch.randint(0, 2 ** bits, (count,), generator=g, dtype=torch.int64)
    elif dist == "skewed":        # most pairs in a handful of tiles, long runs of equal digits
        keys = (torch.randint(0, 2 ** bits, (count,), generator=g, dtype=torch.int64) *
                (torch.rand(count, generator=g) < 0.05)) + 37 * (torch.rand(count, generator=g) < 0.5)
    else:
        keys = torch.full((count,), 200, dtype=torch.int64)
    keys = keys & (2 ** bits - 1)
    vals = torch.arange(count, dtype=torch.int32)
    k32 = (keys & 0xFFFFFFFF).to(torch.int64)
    k32 = torch.where(k32 >= 2 ** 31, k32 - 2 ** 32, k32).to(torch.int32)          # uint32 bits in int32 storage
    ko, vo = utils.sort_pairs32(k32.to(cuda), vals.to(cuda), bits)
    order = _stable_ref(keys, vals)
    assert torch.equal(vo.cpu(), vals[order]), "values (= stability) differ"
    assert torch.equal(ko.cpu(), k32[order])
    # tile-index variant: ids only + bins from the per-tile counts of the last sweep
    if bits <= 16:
        ids, bins = utils.sort_tile_pairs(k32.to(cuda), vals.to(cuda), 2 ** bits)
        assert torch.equal(ids.cpu(), vals[order])
        cnt = torch.bincount(keys, minlength=2 ** bits)
        end = torch.cumsum(cnt, 0)
        ref_bins = torch.stack([end - cnt, end], dim=1)
        ref_bins[cnt == 0] = 0
        assert torch.equal(bins.cpu().long(), ref_bins)


@pytest.mark.parametrize("count", [1, 33, TILE64 - 1, TILE64, TILE64 + 1, 2 * TILE64 + 5, 700_001])
@pytest.mark.parametrize("begin_bit,end_bit", [(0, 42), (32, 45), (0, 64), (0, 37), (5, 6)])
def test_sort_pairs64_matches_stable_sort(cuda, sb_lib, count, begin_bit, end_bit):
    g = torch.Generator(device="cpu").manual_seed(count + 7 * end_bit)
    hi = torch.randint(0, 2 ** 31, (count,), generator=g, dtype=torch.int64)
    lo = torch.randint(0, 2 ** 32, (count,), generator=g, dtype=torch.int64)
    keys = (hi << 32) | lo
    keys[::3] = keys[0]                                          # many exact ties
    vals = torch.arange(count, dtype=torch.int32)
    lib = sb_lib
    kd, vd = keys.to(cuda), vals.to(cuda)
    ko, vo = torch.empty_like(kd), torch.empty_like(vd)
    tmp = torch.empty(lib.sb_sort_temp_bytes(count), dtype=torch.uint8, device=cuda)
    _abi.check(lib.sb_sort_pairs_range(count, begin_bit, end_bit, _abi.ptr(kd, torch.int64, "k"),
                                       _abi.ptr(vd, torch.int32, "v"), _abi.ptr(ko, torch.int64, "ko"),
                                       _abi.ptr(vo, torch.int32, "vo"), _abi.ptr(tmp, torch.uint8, "t"), tmp.numel(),
                                       _abi.stream()), "sb_sort_pairs_range")
    width = end_bit - begin_bit
    field = (keys >> begin_bit) & ((1 << width) - 1) if width < 64 else keys
    if width == 64:       # all 64 bits as UNSIGNED: flip the sign bit for the signed reference sort
        field = keys ^ (-2 ** 63)
    order = _stable_ref(field, vals)
    assert torch.equal(vo.cpu(), vals[order])
    assert torch.equal(ko.cpu(), keys[order])
    assert torch.equal(kd.cpu(), keys), "the input must not be written"


@pytest.mark.parametrize("count", [1, 15, 16, 4095, 4096, 4097, 8 * 4096 + 3, 3_200_000])
def test_cumsum_matches_torch(cuda, sb_lib, count):
    g = torch.Generator(device="cpu").manual_seed(count)
    x = torch.randint(0, 50, (count,), generator=g, dtype=torch.int32)
    total, cum = utils.compute_cumulative_intersects(x.to(cuda))
    ref = torch.cumsum(x, 0, dtype=torch.int32)
    assert torch.equal(cum.cpu(), ref) and total == int(ref[-1])
    # an unaligned view (the scan's vector path must fall back)
    y = torch.zeros(count + 1, dtype=torch.int32, device=cuda)
    y[1:] = x.to(cuda)
    total2, cum2 = utils.compute_cumulative_intersects(y[1:])
    assert torch.equal(cum2.cpu(), ref)


def test_large_sort_against_cub_reference(cuda, sb_lib):
    """20M pairs of tile-shaped keys: identical to cub::DeviceRadixSort (test-only build)."""
    from cub_ref import build as cub_build
    cub = cub_build.load()
    count, bits = 20_000_003, 16
    g = torch.Generator(device=cuda).manual_seed(3)
    keys = torch.randint(0, 65280, (count,), generator=g, device=cuda, dtype=torch.int32)
    vals = torch.randint(0, 1_000_000, (count,), generator=g, device=cuda, dtype=torch.int32)
    ko, vo = utils.sort_pairs32(keys, vals, bits)
    rk, rv = torch.empty_like(keys), torch.empty_like(vals)
    tmp = torch.empty(cub.cubref_sort32_temp(count), dtype=torch.uint8, device=cuda)
    st = _abi.stream()
    rc = cub.cubref_sort32(count, 0, bits, keys.data_ptr(), vals.data_ptr(), rk.data_ptr(), rv.data_ptr(),
                           tmp.data_ptr(), tmp.numel(), st)
    assert rc == 0
    assert torch.equal(ko, rk) and torch.equal(vo, rv)
    ids, bins = utils.sort_tile_pairs(keys, vals, 65280)
    assert torch.equal(ids, rv)
    cnt = torch.bincount(keys.long(), minlength=65280)
    assert torch.equal((bins[:, 1] - bins[:, 0]).long(), cnt)
