"""Host logic of the fused path that needs no GPU: the per-call-site RenderPlan (segment capacity and rect-list
pool learnt from the totals the device reports) and the flat gradient layout the view-sharded all-reduce relies on."""
import types

from splatformer_b200 import render


class _Lib:
    def sb_tile_sort_capacity(self):
        return 8192


def _totals(plan, key, value):
    return render._Totals(None, None, key, plan, value=value)


def test_plan_learns_capacity_pool_and_global_sort_per_call_site():
    lib = _Lib()
    a, b = render.RenderPlan(), render.RenderPlan(tile_capacity=64)
    key = (32, 1024, 100_000)
    assert a.capacity(lib, *key) == render.DEFAULT_TILE_CAPACITY and b.capacity(lib, *key) == 64
    # before anything is known: 12 ints per (view, Gaussian), never more than the fixed-capacity layout
    assert a.pool_capacity(key, 2048) == (12 * 32 * 100_000 + (1 << 20) - 1) >> 20 << 20
    assert b.pool_capacity((2, 36, 3000), 64) == 2 * 36 * 8 * 64
    # a pass reports 7.6 M intersections, no overflow, 12.2 M rect-list entries
    _totals(a, key, (7_575_408, 0, 12_208_752)).read()
    assert a.capacity(lib, *key) == render.DEFAULT_TILE_CAPACITY
    assert 12_208_752 * 1.3 <= a.pool_capacity(key, 2048) <= 12_208_752 * 1.3 + (2 << 20)
    # tile overflow doubles the capacity of THAT plan only; pool overflow alone does not
    _totals(b, key, (1000, 1, 500)).read()
    _totals(b, key, (1000, 2, 900)).read()
    assert b.capacity(lib, *key) == 128 and b.pool_used[key] == 900 and not a.overflowed
    # a shape that averages more than 16 tiles per Gaussian goes to the global sort from then on
    big = (8, 8160, 1_000_000)
    _totals(a, big, (47_000_000 * 8, 1, 0)).read()
    assert a.capacity(lib, *big) == 0
    # capacities past the library's limit / the 2^31 index space fall back to the global sort as well
    c = render.RenderPlan(tile_capacity=1 << 14)
    assert c.capacity(lib, *key) == 0
    assert render.RenderPlan().capacity(lib, 64, 32400, 10) == 0          # 64 * 32400 * 2048 >= 2^31


def test_totals_are_folded_once_and_polled_without_blocking():
    plan = render.RenderPlan()
    key = (2, 36, 3000)
    t = _totals(plan, key, (10, 1, 5))
    plan.pending.append(t)
    plan.poll()
    assert plan.pending == [] and plan.overflowed[key] == 1
    t.read(); t.read()
    assert plan.overflowed[key] == 1                                      # folded once
    pending = types.SimpleNamespace(ready=lambda: False)
    plan.pending.append(pending)
    plan.poll()
    assert plan.pending == [pending]                                       # not there yet: left alone, no wait


def test_grad_layout_blocks_are_disjoint_aligned_and_cover_all_parameters():
    n, k = 1001, 16
    blocks, size = render.grad_layout(n, (n, 3), (n, k - 1, 3))
    names = [b[0] for b in blocks]
    assert names == ["means", "scales", "quats", "opacities", "color_src", "color_rest"]
    end = 0
    for name, off, shape in blocks:
        cnt = 1
        for d in shape:
            cnt *= d
        assert off % 4 == 0 and off >= end and off - end <= 3, name        # 16-byte aligned, no real holes
        end = off + cnt
    assert end <= size <= end + 3
    assert sum(b[2][0] * (b[2][1] if len(b[2]) == 2 else b[2][1] * b[2][2]) for b in blocks) == n * (11 + 3 * k)
    blocks2, _ = render.grad_layout(n, (n, 3), None)                        # degree-0 path: no features_rest
    assert [b[0] for b in blocks2][-1] == "color_src"
