"""Parity of the CUDA path (through the gsplat-compatible API / C ABI) against the oracle.
Runs on the B200 box: python -m pytest tests -m gpu."""
import math

import pytest
import torch

import gsplat
from oracle import ref_torch as R
from splatformer_b200 import synth
import helpers as H

pytestmark = pytest.mark.gpu

CASES = [
    # n, sh degree, image size (W,H), block width, scale multiplier
    (600, 2, (64, 64), 16, 0.6),
    (4096, 0, (128, 128), 16, 0.15),
    (3000, 3, (200, 120), 16, 0.3),     # ragged: neither side a multiple of 16
    (2000, 1, (96, 80), 12, 0.3),       # block width != 16
]


def _scene(n, deg, size, mult, seed=11, views=2):
    gs = synth.shell_scene(n, deg, seed=seed, scale_mult=mult)
    cams = synth.orbit_cameras(views, size[0], size[1], background=(0.1, 0.2, 0.3))
    return gs, cams


@pytest.mark.parametrize("n,deg,size,bw,mult", CASES)
def test_sh_forward_backward(cuda, sb_lib, n, deg, size, bw, mult):
    gs, cams = _scene(n, max(deg, 1), size, mult)
    inp = H.view_inputs(gs, cams, 0)
    for use in range(0, max(deg, 1) + 1):
        c_ref = inp["coeffs"].clone().requires_grad_()
        out_ref = R.spherical_harmonics(use, inp["viewdirs"], c_ref)
        g = torch.randn_like(out_ref)
        out_ref.backward(g)
        c = inp["coeffs"].to(cuda).requires_grad_()
        out = gsplat.spherical_harmonics(use, inp["viewdirs"].to(cuda), c)
        out.backward(g.to(cuda))
        assert torch.allclose(out.cpu(), out_ref, atol=1e-5, rtol=1e-5)
        assert torch.allclose(c.grad.cpu(), c_ref.grad, atol=1e-5, rtol=1e-5)


@pytest.mark.parametrize("n,deg,size,bw,mult", CASES)
def test_projection_bit_exact_and_grads(cuda, sb_lib, n, deg, size, bw, mult):
    gs, cams = _scene(n, deg, size, mult)
    k = H.cam_scalars(cams)
    for view in range(2):
        inp = H.view_inputs(gs, cams, view)
        leaves = [inp[x].clone().requires_grad_() for x in ("means", "scales", "quats")]
        ref = R.project_gaussians(leaves[0], leaves[1], 1, leaves[2], inp["viewmat"], k["fx"], k["fy"],
                                  k["cx"], k["cy"], k["H"], k["W"], bw)
        dl = [inp[x].to(cuda).requires_grad_() for x in ("means", "scales", "quats")]
        got = gsplat.project_gaussians(dl[0], dl[1], 1, dl[2], inp["viewmat"].to(cuda), k["fx"], k["fy"],
                                       k["cx"], k["cy"], k["H"], k["W"], bw)
        names = ["xys", "depths", "radii", "conics", "compensation", "num_tiles_hit", "cov3d"]
        assert got[2].dtype == torch.int32 and got[5].dtype == torch.int32
        for nm, a, b in zip(names, got, ref):
            assert torch.equal(a.detach().cpu(), b.detach()), f"{nm} is not bit-identical to the oracle"
        assert int((ref[2] > 0).sum()) > n // 10
        torch.manual_seed(view)
        ws = [torch.randn_like(ref[i]) for i in (0, 1, 3, 4, 6)]
        loss_ref = sum((ref[i] * w).sum() for i, w in zip((0, 1, 3, 4, 6), ws))
        loss_ref.backward()
        loss = sum((got[i] * w.to(cuda)).sum() for i, w in zip((0, 1, 3, 4, 6), ws))
        loss.backward()
        for nm, a, b in zip(("v_means", "v_scales", "v_quats"), dl, leaves):
            H.assert_grad_close(a.grad, b.grad, nm)


def test_projection_viewmat_grad(cuda, sb_lib):
    gs, cams = _scene(500, 0, (64, 64), 0.6)
    k = H.cam_scalars(cams)
    inp = H.view_inputs(gs, cams, 1)
    vm_ref = inp["viewmat"].clone().requires_grad_()
    ref = R.project_gaussians(inp["means"], inp["scales"], 1, inp["quats"], vm_ref, k["fx"], k["fy"], k["cx"],
                              k["cy"], k["H"], k["W"], 16)
    w0, w3 = torch.randn_like(ref[0]), torch.randn_like(ref[3])
    ((ref[0] * w0).sum() + (ref[3] * w3).sum() + ref[1].sum()).backward()
    vm = inp["viewmat"].to(cuda).requires_grad_()
    got = gsplat.project_gaussians(inp["means"].to(cuda), inp["scales"].to(cuda), 1, inp["quats"].to(cuda), vm,
                                   k["fx"], k["fy"], k["cx"], k["cy"], k["H"], k["W"], 16)
    ((got[0] * w0.to(cuda)).sum() + (got[3] * w3.to(cuda)).sum() + got[1].sum()).backward()
    H.assert_grad_close(vm.grad, vm_ref.grad, "v_viewmat")


@pytest.mark.parametrize("n,deg,size,bw,mult", CASES)
def test_tiling_bit_exact(cuda, sb_lib, n, deg, size, bw, mult):
    gs, cams = _scene(n, deg, size, mult)
    k = H.cam_scalars(cams)
    inp = H.view_inputs(gs, cams, 0)
    xys, depths, radii, conics, comp, tiles, cov3d = R.project_gaussians(
        inp["means"], inp["scales"], 1, inp["quats"], inp["viewmat"], k["fx"], k["fy"], k["cx"], k["cy"],
        k["H"], k["W"], bw)
    tb = ((k["W"] + bw - 1) // bw, (k["H"] + bw - 1) // bw, 1)
    ni_ref, cum_ref = R.compute_cumulative_intersects(tiles)
    ref = R.bin_and_sort_gaussians(n, ni_ref, xys, depths, radii, cum_ref, tb, bw)
    ni, cum = gsplat.compute_cumulative_intersects(tiles.to(cuda))
    assert ni == ni_ref and torch.equal(cum.cpu(), cum_ref)
    got = gsplat.bin_and_sort_gaussians(n, ni, xys.to(cuda), depths.to(cuda), radii.to(cuda), cum, tb, bw)
    for nm, a, b in zip(("isect_ids", "gaussian_ids", "isect_ids_sorted", "gaussian_ids_sorted", "tile_bins"),
                        got, ref):
        assert a.dtype == b.dtype, nm
        assert torch.equal(a.cpu(), b), f"{nm} differs from the oracle"
    srt = got[2].cpu()
    assert bool((srt[1:] >= srt[:-1]).all())


def _raster_both(cuda, gs, cams, view, deg, bw, dtype_ref=torch.float64):
    """Rasterise the same float32 projected splats with the CUDA path and the float64 oracle."""
    k = H.cam_scalars(cams)
    inp = H.view_inputs(gs, cams, view)
    with torch.no_grad():
        xys, depths, radii, conics, comp, tiles, cov3d = R.project_gaussians(
            inp["means"], inp["scales"], 1, inp["quats"], inp["viewmat"], k["fx"], k["fy"], k["cx"], k["cy"],
            k["H"], k["W"], bw)
        colors = H.colors_from(R, inp, deg)
    ref_leaves = [t.to(dtype_ref).clone().requires_grad_() for t in (xys, conics, colors, inp["opacities"])]
    img_ref, alpha_ref, aux = R.rasterize_gaussians(
        ref_leaves[0], depths, radii, ref_leaves[1], tiles, ref_leaves[2], ref_leaves[3], k["H"], k["W"], bw,
        background=inp["background"].to(dtype_ref), return_alpha=True, return_aux=True)
    leaves = [t.to(cuda).clone().requires_grad_() for t in (xys, conics, colors, inp["opacities"])]
    img, alpha = gsplat.rasterize_gaussians(leaves[0], depths.to(cuda), radii.to(cuda), leaves[1], tiles.to(cuda),
                                            leaves[2], leaves[3], k["H"], k["W"], bw,
                                            background=inp["background"].to(cuda), return_alpha=True)
    return (img, alpha, leaves), (img_ref, alpha_ref, ref_leaves, aux)


@pytest.mark.parametrize("n,deg,size,bw,mult", CASES)
def test_rasterize_forward_and_backward(cuda, sb_lib, n, deg, size, bw, mult):
    gs, cams = _scene(n, deg, size, mult)
    for view in range(2):
        (img, alpha, leaves), (img_ref, alpha_ref, ref_leaves, aux) = _raster_both(cuda, gs, cams, view, deg, bw)
        solid = aux["margin"] > H.FRAGILE_MARGIN
        assert float((~solid).float().mean()) < 2e-3, "too many fragile pixels for a meaningful test"
        assert float(alpha_ref.mean()) > 0.02, "scene does not cover the image"
        err_img = (img.cpu().double() - img_ref.detach()).abs().amax(dim=-1)
        err_a = (alpha.cpu().double() - alpha_ref.detach()).abs()
        assert float(err_img[solid].max()) <= H.IMG_ATOL, float(err_img[solid].max())
        assert float(err_a[solid].max()) <= H.IMG_ATOL
        # fragile pixels can flip one decision: bounded by one alpha step
        assert float(err_img.max()) <= 1.5 / 255 + 1e-3
        # backward: random cotangents, zeroed on the fragile pixels
        torch.manual_seed(3 + view)
        v_img = torch.randn(img_ref.shape, dtype=torch.float64) * solid[..., None]
        v_alpha = torch.randn(alpha_ref.shape, dtype=torch.float64) * solid
        ((img_ref * v_img).sum() + (alpha_ref * v_alpha).sum()).backward()
        ((img * v_img.float().to(cuda)).sum() + (alpha * v_alpha.float().to(cuda)).sum()).backward()
        for nm, a, b in zip(("v_xys", "v_conics", "v_colors", "v_opacity"), leaves, ref_leaves):
            H.assert_grad_close(a.grad, b.grad, nm)


def test_rasterize_empty_scene_and_no_alpha(cuda, sb_lib):
    n = 64
    z = lambda *s, dt=torch.float32: torch.zeros(*s, dtype=dt, device=cuda)
    bg = torch.tensor([0.2, 0.4, 0.6], device=cuda)
    img, alpha = gsplat.rasterize_gaussians(z(n, 2), z(n), z(n, dt=torch.int32), z(n, 3), z(n, dt=torch.int32),
                                            z(n, 3), z(n, 1), 40, 56, 16, background=bg, return_alpha=True)
    assert img.shape == (40, 56, 3) and alpha.shape == (40, 56)
    assert torch.allclose(img, bg.expand(40, 56, 3)) and bool((alpha == 1).all())   # upstream quirk (A.5)
    only = gsplat.rasterize_gaussians(z(n, 2), z(n), z(n, dt=torch.int32), z(n, 3), z(n, dt=torch.int32),
                                      z(n, 3), z(n, 1), 40, 56, 16)
    assert isinstance(only, torch.Tensor) and bool((only == 1).all())               # default background = ones


def test_api_validation_errors(cuda, sb_lib):
    n = 8
    m = torch.rand(n, 3, device=cuda); s = torch.rand(n, 3, device=cuda) * 0.1
    q = torch.randn(n, 4, device=cuda); qn = q / q.norm(dim=-1, keepdim=True)
    vm = torch.eye(4, device=cuda)[:3]
    with pytest.raises(AssertionError):
        gsplat.project_gaussians(m, s, 1, qn, vm, 100., 100., 32., 32., 64, 64, 17)
    with pytest.raises(AssertionError):
        gsplat.project_gaussians(m, s, 1, q * 3, vm, 100., 100., 32., 32., 64, 64, 16)
    out = gsplat.project_gaussians(m, s, 1, qn, torch.eye(4, device=cuda), 100., 100., 32., 32., 64, 64, 16)
    assert len(out) == 7
    with pytest.raises(ValueError):
        gsplat.rasterize_gaussians(torch.zeros(n, 3, device=cuda), out[1], out[2], out[3], out[5],
                                   torch.zeros(n, 3, device=cuda), torch.zeros(n, 1, device=cuda), 64, 64, 16)
    with pytest.raises(AssertionError):
        gsplat.rasterize_gaussians(out[0], out[1], out[2], out[3], out[5], torch.zeros(n, 3, device=cuda),
                                   torch.zeros(n, 1, device=cuda), 64, 64, 16, background=torch.zeros(4, device=cuda))
    with pytest.raises(RuntimeError):
        gsplat.spherical_harmonics(0, torch.zeros(n, 3), torch.zeros(n, 1, 3))   # CPU tensors: no CPU path


def test_full_view_pipeline_matches_oracle(cuda, sb_lib):
    """rasterize_gaussians_to_singleimg logic (gs_utils.py:29-114) end to end, fwd + grads to raw params."""
    gs, cams = _scene(3000, 3, (160, 128), 0.3, seed=5)
    k = H.cam_scalars(cams)
    gt = synth.target_images(2, k["H"], k["W"], seed=9)
    raw_ref = {n_: v.clone().requires_grad_() for n_, v in gs.items()}
    raw = {n_: v.to(cuda).requires_grad_() for n_, v in gs.items()}
    loss_ref, loss, masks = 0.0, 0.0, []
    for view in range(2):
        rgb_ref, a_ref, aux = R.render_view(raw_ref, cams["camera_to_worlds"][view], cams["cx"], cams["cy"],
                                            cams["fx"], cams["fy"], cams["width"], cams["height"],
                                            cams["background_color"], return_aux=True)
        solid = (aux["margin"] > H.FRAGILE_MARGIN)
        # same wrapper logic on the CUDA backend
        c2w = cams["camera_to_worlds"][view].to(cuda)
        vm = R.viewmat_from_c2w(cams["camera_to_worlds"][view])[:3].to(cuda)
        scales = torch.exp(raw["scales"]); quats = raw["quats"] / raw["quats"].norm(dim=-1, keepdim=True)
        opac = torch.sigmoid(raw["opacities"])
        coeffs = torch.cat([raw["features_dc"].unsqueeze(1), raw["features_rest"]], dim=1)
        dirs = raw["means"].detach() - c2w[:3, 3]; dirs = dirs / dirs.norm(dim=-1, keepdim=True)
        rgbs = torch.clamp(gsplat.spherical_harmonics(3, dirs, coeffs) + 0.5, min=0.0)
        xys, depths, radii, conics, comp, tiles, cov3d = gsplat.project_gaussians(
            raw["means"], scales, 1, quats, vm, k["fx"], k["fy"], k["cx"], k["cy"], k["H"], k["W"], 16)
        rgb, alpha = gsplat.rasterize_gaussians(xys, depths, radii, conics, tiles, rgbs, opac, k["H"], k["W"], 16,
                                                background=cams["background_color"].to(cuda), return_alpha=True)
        rgb = torch.clamp(rgb, max=1.0)
        err = (rgb.detach().cpu() - rgb_ref.detach()).abs().amax(-1)
        assert float(err[solid].max()) <= H.IMG_ATOL
        mask = solid[..., None].float()
        masks.append(solid)
        loss_ref = loss_ref + ((rgb_ref - gt[view]).abs() * mask).mean()
        loss = loss + ((rgb - gt[view].to(cuda)).abs() * mask.to(cuda)).mean()
    loss.backward()
    assert abs(float(loss) - float(loss_ref)) < 1e-5
    ref, _ = H.oracle_f64_grads(gs, cams, gt, (0, 1), masks)            # float64 oracle, same masked loss
    for name in gs:
        H.assert_grad_close(raw[name].grad, ref[name], name, rtol=H.GRAD_RTOL)


@pytest.mark.parametrize("channels", [1, 2, 4, 7])
def test_rasterize_n_channels(cuda, sb_lib, channels):
    """Colour counts other than 3 (upstream's nd_rasterize path): image, alpha and all
    four gradients against the float64 oracle."""
    gs, cams = _scene(1500, 0, (88, 72), 0.3, seed=5)
    k = H.cam_scalars(cams)
    inp = H.view_inputs(gs, cams, 1)
    with torch.no_grad():
        xys, depths, radii, conics, comp, tiles, cov3d = R.project_gaussians(
            inp["means"], inp["scales"], 1, inp["quats"], inp["viewmat"], k["fx"], k["fy"], k["cx"], k["cy"],
            k["H"], k["W"], 16)
    gen = torch.Generator().manual_seed(channels)
    colors = torch.rand(xys.shape[0], channels, generator=gen)
    bg = torch.rand(channels, generator=gen)
    ref_leaves = [t.double().clone().requires_grad_() for t in (xys, conics, colors, inp["opacities"])]
    img_ref, alpha_ref, aux = R.rasterize_gaussians(
        ref_leaves[0], depths, radii, ref_leaves[1], tiles, ref_leaves[2], ref_leaves[3], k["H"], k["W"], 16,
        background=bg.double(), return_alpha=True, return_aux=True)
    leaves = [t.to(cuda).clone().requires_grad_() for t in (xys, conics, colors, inp["opacities"])]
    img, alpha = gsplat.rasterize_gaussians(leaves[0], depths.to(cuda), radii.to(cuda), leaves[1], tiles.to(cuda),
                                            leaves[2], leaves[3], k["H"], k["W"], 16, background=bg.to(cuda),
                                            return_alpha=True)
    assert img.shape == (k["H"], k["W"], channels) and alpha.shape == (k["H"], k["W"])
    solid = aux["margin"] > H.FRAGILE_MARGIN
    assert float((img.cpu().double() - img_ref.detach()).abs().amax(-1)[solid].max()) <= H.IMG_ATOL
    assert float((alpha.cpu().double() - alpha_ref.detach()).abs()[solid].max()) <= H.IMG_ATOL
    v_img = torch.randn(img_ref.shape, dtype=torch.float64, generator=gen) * solid[..., None]
    v_alpha = torch.randn(alpha_ref.shape, dtype=torch.float64, generator=gen) * solid
    ((img_ref * v_img).sum() + (alpha_ref * v_alpha).sum()).backward()
    ((img * v_img.float().to(cuda)).sum() + (alpha * v_alpha.float().to(cuda)).sum()).backward()
    for nm, a, b in zip(("v_xys", "v_conics", "v_colors", "v_opacity"), leaves, ref_leaves):
        H.assert_grad_close(a.grad, b.grad, nm)
    with pytest.raises(AssertionError):
        gsplat.rasterize_gaussians(leaves[0], depths.to(cuda), radii.to(cuda), leaves[1], tiles.to(cuda), leaves[2],
                                   leaves[3], k["H"], k["W"], 16, background=torch.ones(channels + 1, device=cuda))


def test_compute_cov2d_bounds(cuda, sb_lib):
    gen = torch.Generator().manual_seed(0)
    l = torch.randn(4000, 2, 2, generator=gen) * torch.logspace(-1, 1.5, 4000)[:, None, None]
    cov = l @ l.transpose(1, 2) + 0.3 * torch.eye(2)
    cov2d = torch.stack([cov[:, 0, 0], cov[:, 0, 1], cov[:, 1, 1]], dim=1).contiguous()
    cov2d[:5] = torch.tensor([[1.0, 1.0, 1.0], [0.0, 0.0, 0.0], [4.0, 2.0, 1.0], [1.0, 0.0, 1.0], [2.0, 0.0, 0.5]])
    conic_ref, radius_ref = R.compute_cov2d_bounds(cov2d)
    conic, radius = gsplat.utils.compute_cov2d_bounds(cov2d.to(cuda))
    assert radius.dtype == torch.int32 and conic.shape == (4000, 3)
    assert bool((radius[:3].cpu() == 0).all()) and bool((conic[:3] == 0).all())     # singular: det == 0
    assert int(radius[3]) == 4 and torch.allclose(conic[3].cpu(), torch.tensor([1.0, 0.0, 1.0]))
    # the kernel rounds once per operation like the oracle's float32 chain: identical bits
    assert torch.equal(conic.cpu(), conic_ref)
    assert torch.equal(radius.cpu(), radius_ref)


@pytest.mark.parametrize("n,size,views,mult", [(3000, (200, 120), 3, 0.3), (800, (64, 48), 2, 2.5)])
def test_depth_ordered_binning_equals_full_key_sort(cuda, sb_lib, n, size, views, mult):
    """The fused path's global sort (depth pre-order + emission + radix sort over the tile bits only) gives
    the same sorted keys, ids and bins as sorting the whole 64-bit key (= the oracle's order: ascending
    depth bits per tile, ties by ascending id)."""
    from splatformer_b200 import utils as U
    gs, cams = _scene(n, 0, size, mult, views=views)
    gs["means"][: n // 40] = gs["means"][0]                     # exact depth ties
    k = H.cam_scalars(cams)
    a = H.activated(gs, cuda)
    tb = ((k["W"] + 15) // 16, (k["H"] + 15) // 16, 1)
    outs = []
    for v in range(views):
        vm = R.viewmat_from_c2w(cams["camera_to_worlds"][v])[:3].contiguous().to(cuda)
        outs.append(gsplat.project_gaussians(a["means"], a["scales"], 1, a["quats"], vm, k["fx"], k["fy"], k["cx"],
                                             k["cy"], k["H"], k["W"], 16))
    xys = torch.stack([o[0] for o in outs]).contiguous()
    depths = torch.stack([o[1] for o in outs]).contiguous()
    radii = torch.stack([o[2] for o in outs]).contiguous()
    tiles = torch.stack([o[5] for o in outs]).contiguous()
    ni, cum = U.compute_cumulative_intersects(tiles)
    _, _, keys_ref, ids_ref, bins_ref = U.bin_and_sort_gaussians(n, ni, xys.view(-1, 2), depths.view(-1),
                                                                 radii.view(-1), cum, tb, 16, num_views=views)
    ni2, keys, ids, bins, view_ends = U.bin_depth_ordered(views, n, xys.view(-1, 2), 2, depths.view(-1),
                                                          radii.view(-1), tiles, tb, 16)
    assert ni2 == ni and ni > 2 * n
    assert torch.equal(keys, keys_ref) and torch.equal(ids, ids_ref) and torch.equal(bins, bins_ref)
    assert torch.equal(view_ends.cpu(), cum.view(views, n)[:, -1].cpu())
    # compact variant (uint32 tile index as the key): same ids and bins, keys = the tile part
    ni3, tiles_sorted, ids3, bins3, _ = U.bin_depth_ordered(views, n, xys.view(-1, 2), 2, depths.view(-1),
                                                            radii.view(-1), tiles, tb, 16, compact=True)
    assert ni3 == ni and torch.equal(ids3, ids_ref) and torch.equal(bins3, bins_ref)
    assert tiles_sorted is None          # the tile-index sort keeps only the ids; the bins carry the tiles
