"""The hand-written radix sort (csrc/radix_sort.cu) against cub::DeviceRadixSort (test-only build,
tests/cub_ref) on the key shapes of BASELINE configs 4 and 5, on this box:
  * pure pair sort, uint32 keys of 16 significant bits (what the fused path's global sort does);
  * the tile-index pipeline: ours = sort keeping only the ids + per-tile counts + bins; library = CUB pair
    sort + bin-edge kernel over the sorted keys (what round 1 shipped);
  * 64-bit (tile | depth) keys, 42 bits (the gsplat-API path), 47 M pairs.
CUDA events, 5 warm-ups, median of 10; prints one JSON line."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from splatformer_b200 import _abi, utils  # noqa: E402
from cub_ref import build as cub_build  # noqa: E402

dev = torch.device("cuda:0")
torch.cuda.set_device(dev)
lib = _abi.load()
cub = cub_build.load()


def timed(fn, warm=5, reps=10):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


def tile_like_keys(count, tiles, seed):
    """Depth-ordered emission: runs of consecutive tile indices (a Gaussian's rows), random starts."""
    g = torch.Generator(device=dev).manual_seed(seed)
    start = torch.randint(0, tiles, (count // 6 + 1,), generator=g, device=dev, dtype=torch.int32)
    keys = (start[:, None] + torch.arange(6, device=dev, dtype=torch.int32)[None, :]).reshape(-1)[:count] % tiles
    vals = torch.randint(0, 6_000_000, (count,), generator=g, device=dev, dtype=torch.int32)
    return keys.contiguous(), vals


out = {}
ONLY = os.environ.get("SORT_BENCH_ONLY", "")
CASES = (("config4_47M", 47_000_000, 65280), ("config5_420M", 420_000_000, 64800))
for name, count, tiles in [c for c in CASES if not ONLY or c[0] == ONLY]:
    keys, vals = tile_like_keys(count, tiles, 1)
    bits = (tiles - 1).bit_length()
    ko, vo = torch.empty_like(keys), torch.empty_like(vals)
    tmp = torch.empty(lib.sb_sort_tiles_global_temp_bytes(count, tiles), dtype=torch.uint8, device=dev)
    ctmp = torch.empty(cub.cubref_sort32_temp(count), dtype=torch.uint8, device=dev)
    counts = torch.zeros(tiles, dtype=torch.int32, device=dev)
    bins = torch.zeros(tiles, 2, dtype=torch.int32, device=dev)
    st = _abi.stream()
    P, I32 = _abi.ptr, torch.int32

    def ours_pairs():
        _abi.check(lib.sb_sort_pairs32(count, bits, P(keys, I32, "k"), P(vals, I32, "v"), P(ko, I32, "ko"), P(vo, I32, "vo"),
                                       P(tmp, torch.uint8, "t"), tmp.numel(), st), "sort32")

    def cub_pairs():
        assert cub.cubref_sort32(count, 0, bits, keys.data_ptr(), vals.data_ptr(), ko.data_ptr(), vo.data_ptr(),
                                 ctmp.data_ptr(), ctmp.numel(), st) == 0

    def ours_pipeline():
        _abi.check(lib.sb_sort_tiles_global(count, tiles, P(keys, I32, "k"), P(vals, I32, "v"), P(vo, I32, "vo"),
                                            P(bins, I32, "b"), None, P(tmp, torch.uint8, "t"), tmp.numel(), st), "tiles")

    def cub_pipeline():
        cub_pairs()
        bins.zero_()
        _abi.check(lib.sb_tile_bin_edges32(count, P(ko, I32, "ko"), tiles, P(bins, I32, "b"), st), "edges")

    r = {"pairs": count, "key_bits": bits,
         "ours_pair_sort_ms": timed(ours_pairs), "cub_pair_sort_ms": timed(cub_pairs),
         "ours_ids_plus_bins_ms": timed(ours_pipeline), "cub_sort_plus_bin_edges_ms": timed(cub_pipeline)}
    ours_pipeline(); b1 = bins.clone(); v1 = vo.clone()
    cub_pipeline()
    r["identical"] = bool(torch.equal(b1, bins) and torch.equal(v1, vo))
    r["pair_sort_gbs_ours"] = count * 8 * 2 * ((bits + 7) // 8) / r["ours_pair_sort_ms"] / 1e6
    out[name] = r
    del keys, vals, ko, vo, tmp, ctmp
    torch.cuda.empty_cache()

if ONLY:
    print(json.dumps(out))
    sys.exit(0)
# 64-bit (tile | depth) keys, 42 significant bits
count = 47_000_000
g = torch.Generator(device=dev).manual_seed(2)
keys = (torch.randint(0, 1024, (count,), generator=g, device=dev, dtype=torch.int64) << 32) | \
    torch.randint(0x3f000000, 0x40800000, (count,), generator=g, device=dev, dtype=torch.int64)
vals = torch.arange(count, device=dev, dtype=torch.int32)
ko, vo = torch.empty_like(keys), torch.empty_like(vals)
tmp = torch.empty(lib.sb_sort_temp_bytes(count), dtype=torch.uint8, device=dev)
ctmp = torch.empty(cub.cubref_sort64_temp(count), dtype=torch.uint8, device=dev)
st = _abi.stream()


def ours64():
    _abi.check(lib.sb_sort_pairs(count, 42, _abi.ptr(keys, torch.int64, "k"), _abi.ptr(vals, torch.int32, "v"),
                                 _abi.ptr(ko, torch.int64, "ko"), _abi.ptr(vo, torch.int32, "vo"),
                                 _abi.ptr(tmp, torch.uint8, "t"), tmp.numel(), st), "sort64")


def cub64():
    assert cub.cubref_sort64(count, 0, 42, keys.data_ptr(), vals.data_ptr(), ko.data_ptr(), vo.data_ptr(),
                             ctmp.data_ptr(), ctmp.numel(), st) == 0


r = {"pairs": count, "key_bits": 42, "ours_ms": timed(ours64), "cub_ms": timed(cub64)}
ours64(); v1 = vo.clone(); cub64()
r["identical"] = bool(torch.equal(v1, vo))
out["api_64bit_47M"] = r
print(json.dumps(out))
