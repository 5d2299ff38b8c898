"""The vectorised oracle (masked cumprod + autograd) against the literal per-pixel loops of
oracle/ref_loop.py (sequential walk, hand-derived reverse walk) and against float64 finite
differences.  Two independent statements of SURVEY.md A.5/A.6 must agree."""
import numpy as np
import pytest
import torch

from oracle import ref_loop as L
from oracle import ref_torch as R
from splatformer_b200 import synth
import helpers as H


def _projected(n, size, seed, bw=16, mult=0.6, dtype=torch.float64):
    gs = synth.shell_scene(n, 0, seed=seed, scale_mult=mult)
    cams = synth.orbit_cameras(1, size[0], size[1], background=(0.2, 0.3, 0.1))
    k = H.cam_scalars(cams)
    inp = H.view_inputs(gs, cams, 0)
    with torch.no_grad():
        xys, depths, radii, conics, comp, tiles, cov3d = R.project_gaussians(
            inp["means"], inp["scales"], 1, inp["quats"], inp["viewmat"], k["fx"], k["fy"], k["cx"], k["cy"],
            k["H"], k["W"], bw)
    colors = torch.sigmoid(inp["coeffs"][:, 0, :])
    return (xys.to(dtype), depths, radii, conics.to(dtype), tiles, colors.to(dtype), inp["opacities"].to(dtype),
            inp["background"].to(dtype), k)


@pytest.mark.parametrize("n,size,bw,seed", [(60, (32, 32), 16, 1), (150, (40, 24), 16, 2), (80, (30, 30), 12, 3)])
def test_vectorised_oracle_matches_pixel_loop(n, size, bw, seed):
    xys, depths, radii, conics, tiles, colors, opac, bg, k = _projected(n, size, seed, bw)
    leaves = [t.clone().requires_grad_() for t in (xys, conics, colors, opac)]
    img, alpha, aux = R.rasterize_gaussians(leaves[0], depths, radii, leaves[1], tiles, leaves[2], leaves[3],
                                            k["H"], k["W"], bw, background=bg, return_alpha=True, return_aux=True)
    assert aux["num_intersects"] > n // 2
    torch.manual_seed(seed)
    v_img = torch.randn(img.shape, dtype=torch.float64)
    v_alpha = torch.randn(alpha.shape, dtype=torch.float64)
    ((img * v_img).sum() + (alpha * v_alpha).sum()).backward()
    res = L.rasterize_loop(xys.numpy(), conics.numpy(), colors.numpy(), opac.numpy(),
                           aux["gaussian_ids_sorted"].numpy(), aux["tile_bins"].numpy(), k["H"], k["W"], bw,
                           bg.numpy(), v_out_img=v_img.numpy(), v_out_alpha=v_alpha.numpy())
    assert np.allclose(res["out_img"], img.detach().numpy(), atol=1e-12)
    assert np.allclose(res["out_alpha"], alpha.detach().numpy(), atol=1e-12)
    assert np.array_equal(res["final_idx"], aux["final_idx"].numpy())
    for name, leaf, key in (("v_xy", leaves[0], "v_xy"), ("v_conic", leaves[1], "v_conic"),
                            ("v_colors", leaves[2], "v_colors")):
        assert np.allclose(res[key], leaf.grad.numpy(), rtol=1e-9, atol=1e-12), name
    assert np.allclose(res["v_opacity"], leaves[3].grad.numpy().reshape(-1), rtol=1e-9, atol=1e-12)


def test_projection_gradients_match_finite_differences():
    torch.manual_seed(0)
    n = 6
    gs = synth.shell_scene(n, 0, seed=4, scale_mult=3.0)
    cams = synth.orbit_cameras(1, 64, 48)
    k = H.cam_scalars(cams)
    inp = H.view_inputs(gs, cams, 0, dtype=torch.float64)

    def fn(means, scales, quats, viewmat):
        xys, depths, radii, conics, comp, tiles, cov3d = R.project_gaussians(
            means, scales, 1, quats, viewmat, k["fx"], k["fy"], k["cx"], k["cy"], k["H"], k["W"], 16)
        assert int((radii > 0).sum()) == n
        return xys, depths, conics, comp, cov3d

    args = [inp[x].clone().requires_grad_() for x in ("means", "scales", "quats", "viewmat")]
    assert torch.autograd.gradcheck(fn, args, eps=1e-6, atol=1e-5, rtol=1e-4, nondet_tol=0.0)


def test_raster_gradients_match_finite_differences():
    xys, depths, radii, conics, tiles, colors, opac, bg, k = _projected(25, (32, 32), 9, mult=2.0)
    # keep every decision away from its threshold so that finite differences see a smooth function
    opac = opac.clamp(0.05, 0.6)

    def fn(xy, con, col, op):
        img, alpha, aux = R.rasterize_gaussians(xy, depths, radii, con, tiles, col, op, k["H"], k["W"], 16,
                                                background=bg, return_alpha=True, return_aux=True)
        mask = (aux["margin"] > 1e-3).to(img.dtype)
        return (img * mask[..., None]).sum(dim=(0, 1)), (alpha * mask).sum()

    args = [t.clone().requires_grad_() for t in (xys, conics, colors, opac)]
    assert torch.autograd.gradcheck(fn, args, eps=1e-7, atol=1e-5, rtol=1e-3, nondet_tol=0.0)


def test_wrapper_matches_manual_three_calls():
    gs = synth.shell_scene(400, 2, seed=12, scale_mult=0.6)
    cams = synth.orbit_cameras(2, 48, 40, background=(0.3, 0.2, 0.1))
    k = H.cam_scalars(cams)
    rgbs, alphas = R.render_views(gs, cams)
    assert len(rgbs) == 2 and rgbs[0].shape == (40, 48, 3) and alphas[0].shape == (40, 48, 1)
    inp = H.view_inputs(gs, cams, 1)
    colors = H.colors_from(R, inp, 2)
    xys, depths, radii, conics, comp, tiles, cov3d = R.project_gaussians(
        inp["means"], inp["scales"], 1, inp["quats"], inp["viewmat"], k["fx"], k["fy"], k["cx"], k["cy"], k["H"],
        k["W"], 16)
    img, alpha = R.rasterize_gaussians(xys, depths, radii, conics, tiles, colors, inp["opacities"], k["H"], k["W"],
                                       16, background=inp["background"], return_alpha=True)
    assert torch.allclose(rgbs[1], img.clamp(max=1.0), atol=1e-6)
    assert torch.allclose(alphas[1][..., 0], alpha, atol=1e-6)
    assert float(rgbs[1].max()) <= 1.0


def test_tile_lists_equal_the_global_sort_and_tile_subsets_composite_alike():
    """oracle.tile_lists (tile-by-tile statement used for frames too large for the CPU sort) against
    bin_and_sort_gaussians, and rasterize_forward(tiles=...) against the full frame."""
    xys, depths, radii, conics, tiles_hit, colors, opac, bg, k = _projected(900, (96, 80), 5, 16, mult=0.8,
                                                                           dtype=torch.float32)
    xys[:30] = xys[0]; depths[:30] = depths[0]; radii[:30] = radii[0]            # exact depth ties -> id order
    tb = ((k["W"] + 15) // 16, (k["H"] + 15) // 16, 1)
    from oracle.ref_torch import _tile_bbox
    mnx, mxx, mny, mxy = _tile_bbox(xys[:, 0], xys[:, 1], radii.float(), tb[0], tb[1], 16)
    tiles_hit = torch.where(radii > 0, (mxx - mnx) * (mxy - mny), torch.zeros_like(mnx)).to(torch.int32)
    ni, cum = R.compute_cumulative_intersects(tiles_hit)
    _, _, _, gids_ref, bins_ref = R.bin_and_sort_gaussians(900, ni, xys, depths, radii, cum, tb, 16)
    subset = [0, 3, 7, 11, 17, 22, 29]
    gids, bins = R.tile_lists(xys, depths, radii, tb, 16, subset)
    for t in subset:
        a = gids[int(bins[t, 0]):int(bins[t, 1])]
        b = gids_ref[int(bins_ref[t, 0]):int(bins_ref[t, 1])]
        assert torch.equal(a, b), f"tile {t}"
    full = R.rasterize_forward(xys, conics, colors, opac, gids_ref, bins_ref, k["H"], k["W"], 16, bg)
    part = R.rasterize_forward(xys, conics, colors, opac, gids, bins, k["H"], k["W"], 16, bg, tiles=subset)
    part2 = R.rasterize_forward(xys, conics, colors, opac, gids_ref, bins_ref, k["H"], k["W"], 16, bg, tiles=subset)
    for t in subset:
        ty, tx = divmod(t, tb[0])
        sl = (slice(ty * 16, min(ty * 16 + 16, k["H"])), slice(tx * 16, min(tx * 16 + 16, k["W"])))
        assert torch.equal(full[0][sl], part[0][sl]) and torch.equal(full[1][sl], part[1][sl])
        assert torch.equal(full[0][sl], part2[0][sl])


def test_wrapper_edge_paths_of_the_reference():
    """gs_utils.py:47-51 (a quaternion that does not normalise to unit length becomes (0,0,0,1)) and
    :72-76 (a mean on the camera centre): same image as with the substituted values, finite gradients,
    none for the substituted quaternion."""
    gs = synth.shell_scene(300, 1, seed=3, scale_mult=0.8)
    cams = synth.orbit_cameras(1, 48, 48)
    gs["quats"][5] = 0.0
    gs["quats"][6] = float("nan")
    gs["means"][7] = cams["camera_to_worlds"][0][:3, 3]
    kw = {k: cams[k] for k in ("cx", "cy", "fx", "fy", "width", "height", "background_color")}
    raw = {k: v.clone().requires_grad_() for k, v in gs.items()}
    rgb, _ = R.render_view(raw, cams["camera_to_worlds"][0], **kw)
    gs2 = {k: v.clone() for k, v in gs.items()}
    gs2["quats"][5] = torch.tensor([0.0, 0.0, 0.0, 1.0]); gs2["quats"][6] = torch.tensor([0.0, 0.0, 0.0, 1.0])
    rgb2, _ = R.render_view(gs2, cams["camera_to_worlds"][0], **kw)
    assert torch.isfinite(rgb).all() and torch.equal(rgb, rgb2)
    rgb.sum().backward()
    for name, p in raw.items():
        assert torch.isfinite(p.grad).all(), name
    assert float(raw["quats"].grad[5].abs().sum()) == 0.0 and float(raw["quats"].grad[6].abs().sum()) == 0.0
    assert float(raw["features_rest"].grad[7].abs().sum()) == 0.0          # depth 0: culled, never drawn


def test_projection_and_tiling_loops_match_the_vectorised_oracle_bit_for_bit():
    """Two independently structured statements of A.2-A.4 (unrolled vector expressions vs a per-Gaussian scalar
    loop over 3x3 matrices) must agree in every bit of every output, including the degenerate Gaussians."""
    gs = synth.shell_scene(260, 0, seed=19, scale_mult=1.2)
    gs["means"][:6] = torch.tensor([[0.5, 0.5, 3.0], [5.0, 0.5, 0.5], [0.5, 0.5, 0.5], [0.5001, 0.5, 0.5],
                                    [0.9, 0.1, 0.4], [0.5, -4.0, 0.5]])       # behind / far off-screen / centre
    gs["scales"][6:10] = torch.tensor([[-9.0, -9.0, -9.0], [0.5, -6.0, -6.0], [1.5, 1.5, 1.5], [-3.0, -1.0, -7.0]])
    cams = synth.orbit_cameras(2, 88, 56)
    k = H.cam_scalars(cams)
    for view in range(2):
        inp = H.view_inputs(gs, cams, view)
        with torch.no_grad():
            ref = R.project_gaussians(inp["means"], inp["scales"], 1, inp["quats"], inp["viewmat"], k["fx"], k["fy"],
                                      k["cx"], k["cy"], k["H"], k["W"], 16)
        got = L.project_loop(inp["means"].numpy(), inp["scales"].numpy(), 1.0, inp["quats"].numpy(),
                             inp["viewmat"].numpy(), k["fx"], k["fy"], k["cx"], k["cy"], k["H"], k["W"], 16)
        assert int((ref[2] > 0).sum()) > 100 and int((ref[2] == 0).sum()) >= 2
        for name, a, b in zip(("xys", "depths", "radii", "conics", "compensation", "num_tiles_hit", "cov3d"), got, ref):
            assert np.array_equal(a, b.numpy()), f"view {view}: {name} differs between the two statements"
        tb = ((k["W"] + 15) // 16, (k["H"] + 15) // 16, 1)
        ni, cum = R.compute_cumulative_intersects(ref[5])
        _, _, keys_ref, ids_ref, bins_ref = R.bin_and_sort_gaussians(260, ni, ref[0], ref[1], ref[2], cum, tb, 16)
        keys, ids, bins = L.bin_and_sort_loop(got[0], got[1], got[2], tb, 16)
        assert np.array_equal(keys, keys_ref.numpy()) and np.array_equal(ids, ids_ref.numpy())
        assert np.array_equal(bins, bins_ref.numpy())
