"""The fused multi-view path (splatformer_b200.render / gs_utils) against the per-view
gsplat-compatible API and against the oracle's wrapper."""
import pytest
import torch

import gsplat
from oracle import ref_torch as R
from splatformer_b200 import gs_utils, synth
import helpers as H

pytestmark = pytest.mark.gpu


def _per_view_api(raw, cams, view, deg, cuda):
    """rasterize_gaussians_to_singleimg logic on the per-view API (three gsplat calls)."""
    k = H.cam_scalars(cams)
    c2w = cams["camera_to_worlds"][view].to(cuda)
    vm = gs_utils.viewmats_from_c2w(c2w[None])[0][0]       # same world->camera bits as the fused path
    scales = torch.exp(raw["scales"])
    quats = raw["quats"] / raw["quats"].norm(dim=-1, keepdim=True)
    opac = torch.sigmoid(raw["opacities"])
    if deg == 0:
        rgbs = torch.sigmoid(raw["features_dc"])
    else:
        coeffs = torch.cat([raw["features_dc"].unsqueeze(1), raw["features_rest"]], dim=1)
        dirs = raw["means"].detach() - c2w[:3, 3]
        dirs = dirs / dirs.norm(dim=-1, keepdim=True)
        rgbs = torch.clamp(gsplat.spherical_harmonics(deg, dirs, coeffs) + 0.5, min=0.0)
    xys, depths, radii, conics, comp, tiles, cov3d = gsplat.project_gaussians(
        raw["means"], scales, 1, quats, vm, k["fx"], k["fy"], k["cx"], k["cy"], k["H"], k["W"], 16)
    rgb, alpha = gsplat.rasterize_gaussians(xys, depths, radii, conics, tiles, rgbs, opac, k["H"], k["W"], 16,
                                            background=cams["background_color"].to(cuda), return_alpha=True)
    return torch.clamp(rgb, max=1.0), alpha.unsqueeze(-1)


@pytest.mark.parametrize("n,deg,size,views", [(3000, 3, (160, 128), 3), (2500, 0, (96, 96), 2),
                                              (1500, 1, (72, 100), 4)])
def test_fused_matches_per_view_api(cuda, sb_lib, n, deg, size, views):
    gs = synth.shell_scene(n, deg, seed=21, scale_mult=0.3)
    cams = synth.orbit_cameras(views, size[0], size[1], background=(0.3, 0.1, 0.2))
    gt = synth.target_images(views, size[1], size[0], seed=4).to(cuda)
    raw_a = {k_: v.to(cuda).requires_grad_() for k_, v in gs.items()}
    raw_b = {k_: v.to(cuda).requires_grad_() for k_, v in gs.items()}
    cams_dev = {k_: v.to(cuda) for k_, v in cams.items()}
    rgbs, alphas = gs_utils.rasterize_gaussians_to_multiimgs(raw_a, cams_dev)
    assert len(rgbs) == views and rgbs[0].shape == (size[1], size[0], 3) and alphas[0].shape == (size[1], size[0], 1)
    loss_a = sum((rgbs[v] - gt[v]).abs().mean() + 0.1 * alphas[v].mean() for v in range(views))
    loss_b = 0.0
    for v in range(views):
        rgb, alpha = _per_view_api(raw_b, cams, v, deg, cuda)
        # forward: same device functions underneath; the fused kernels apply exp / sigmoid / the
        # camera inversion themselves (expf vs torch.exp: last-bit differences), so not bit-identical
        assert float((alpha - alphas[v]).abs().max()) <= 1e-5, f"view {v}: fused alpha differs from per-view API"
        assert float((rgb.detach() - rgbs[v].detach()).abs().max()) <= 1e-5
        loss_b = loss_b + (rgb - gt[v]).abs().mean() + 0.1 * alpha.mean()
    loss_a.backward()
    loss_b.backward()
    for name in gs:
        H.assert_grad_close(raw_a[name].grad, raw_b[name].grad, name)   # summation order differs (atomics)


def test_fused_matches_oracle_wrapper(cuda, sb_lib):
    gs = synth.shell_scene(2000, 2, seed=33, scale_mult=0.3)
    cams = synth.orbit_cameras(2, 112, 96, background=(0.0, 0.0, 0.0))
    raw = {k_: v.to(cuda).requires_grad_() for k_, v in gs.items()}
    raw_ref = {k_: v.clone().requires_grad_() for k_, v in gs.items()}
    rgbs, alphas = gs_utils.rasterize_gaussians_to_multiimgs(raw, {k_: v.to(cuda) for k_, v in cams.items()})
    loss, masks = 0.0, []
    gt = synth.target_images(2, 96, 112, seed=8)
    for v in range(2):
        rgb_ref, a_ref, aux = R.render_view(raw_ref, cams["camera_to_worlds"][v], cams["cx"], cams["cy"],
                                            cams["fx"], cams["fy"], cams["width"], cams["height"],
                                            cams["background_color"], return_aux=True)
        solid = aux["margin"] > H.FRAGILE_MARGIN
        err = (rgbs[v].detach().cpu() - rgb_ref.detach()).abs().amax(-1)
        assert float(err[solid].max()) <= H.IMG_ATOL
        assert float((alphas[v].detach().cpu() - a_ref.detach()).abs()[solid].max()) <= H.IMG_ATOL
        m = solid[..., None].float()
        masks.append(solid)
        loss = loss + ((rgbs[v] - gt[v].to(cuda)).abs() * m.to(cuda)).mean()
    loss.backward()
    ref, _ = H.oracle_f64_grads(gs, cams, gt, (0, 1), masks)           # float64 oracle, same masked loss
    for name in gs:
        H.assert_grad_close(raw[name].grad, ref[name], name, rtol=H.GRAD_RTOL)


def test_fused_single_view_and_empty_view(cuda, sb_lib):
    gs = synth.shell_scene(500, 1, seed=2, scale_mult=0.3)
    cams = synth.orbit_cameras(2, 64, 64, background=(0.5, 0.25, 0.125))
    # second camera looks away from the object: nothing intersects -> background, alpha 1 (A.5 quirk)
    away = cams["camera_to_worlds"][1].clone()
    away[:3, :3] = away[:3, :3] @ torch.diag(torch.tensor([-1.0, 1.0, -1.0]))
    cams["camera_to_worlds"][1] = away
    raw = {k_: v.to(cuda) for k_, v in gs.items()}
    cd = {k_: v.to(cuda) for k_, v in cams.items()}
    rgbs, alphas = gs_utils.rasterize_gaussians_to_multiimgs(raw, cd)
    assert torch.allclose(rgbs[1], cd["background_color"].expand(64, 64, 3))
    assert bool((alphas[1] == 1).all())
    assert float(alphas[0].mean()) > 0.05
    rgb0, a0 = gs_utils.rasterize_gaussians_to_singleimg(raw, cd["camera_to_worlds"][0], **{
        k_: cd[k_] for k_ in ("cx", "cy", "fx", "fy", "width", "height", "background_color")})
    assert torch.equal(rgb0, rgbs[0]) and torch.equal(a0, alphas[0])


@pytest.mark.parametrize("n,deg,size,views,opkey", [(2500, 3, (144, 112), 3, "opacities"),
                                                    (1500, 0, (80, 96), 2, "opacities"),
                                                    (1200, 2, (64, 64), 2, "opacities_sigmoid")])
def test_fused_l1_loss_matches_torch_loss(cuda, sb_lib, n, deg, size, views, opkey):
    """render_l1_loss (loss + gradient image inside the raster kernels) == torch L1 on render_scene."""
    gs = synth.shell_scene(n, deg, seed=11, scale_mult=0.3)
    if opkey == "opacities_sigmoid":
        gs["opacities_sigmoid"] = torch.sigmoid(gs.pop("opacities"))
    cams = synth.orbit_cameras(views, size[0], size[1], background=(0.2, 0.4, 0.1))
    gt = synth.target_images(views, size[1], size[0], seed=5).to(cuda)
    cams_dev = {k_: v.to(cuda) for k_, v in cams.items()}
    raw_a = {k_: v.to(cuda).requires_grad_() for k_, v in gs.items()}
    raw_b = {k_: v.to(cuda).requires_grad_() for k_, v in gs.items()}
    weights = torch.linspace(0.5, 1.5, views, device=cuda)            # unequal upstream gradients per view
    l1, mse, imgs = gs_utils.render_l1_loss(raw_a, cams_dev, gt, return_images=True)
    assert l1.shape == (views,) and mse.shape == (views,) and not mse.requires_grad
    rgb, _ = gs_utils.render_scene(raw_b, **cams_dev)
    assert torch.equal(imgs, rgb.detach())
    l1_ref = (rgb - gt).abs().mean(dim=(1, 2, 3))
    mse_ref = ((rgb - gt) ** 2).mean(dim=(1, 2, 3))
    assert torch.allclose(l1, l1_ref, rtol=1e-5, atol=1e-7)
    assert torch.allclose(mse, mse_ref.detach(), rtol=1e-5, atol=1e-7)
    (l1 * weights).sum().backward()
    (l1_ref * weights).sum().backward()
    for name in gs:
        H.assert_grad_close(raw_a[name].grad, raw_b[name].grad, name)


def test_fused_l1_loss_matches_oracle(cuda, sb_lib):
    gs = synth.shell_scene(1800, 1, seed=9, scale_mult=0.3)
    cams = synth.orbit_cameras(2, 96, 80, background=(0.0, 0.0, 0.0))
    gt = synth.target_images(2, 80, 96, seed=12)
    raw = {k_: v.to(cuda).requires_grad_() for k_, v in gs.items()}
    raw_ref = {k_: v.clone().requires_grad_() for k_, v in gs.items()}
    l1, mse = gs_utils.render_l1_loss(raw, {k_: v.to(cuda) for k_, v in cams.items()}, gt.to(cuda))
    rgbs, _ = R.render_views(raw_ref, cams)
    l1_ref = torch.stack([(rgb - g).abs().mean() for rgb, g in zip(rgbs, gt)])
    assert torch.allclose(l1.detach().cpu(), l1_ref.detach(), rtol=1e-4, atol=1e-6)
    l1.sum().backward()
    # gradients against the float64 oracle; the fused L1 cannot set fragile pixels aside, so the elementwise band
    # is checked on the relative L2 norm per tensor (a flipped 1/255 decision moves single entries, not the norm)
    ref, _ = H.oracle_f64_grads(gs, cams, gt, (0, 1))
    for name in gs:
        assert H.rel_l2(raw[name].grad, ref[name]) <= H.GRAD_RTOL, (name, H.rel_l2(raw[name].grad, ref[name]))


def test_fused_l1_loss_uint8_host_targets(cuda, sb_lib):
    """8-bit targets in pinned host memory (uploaded on a side stream) == float targets on the device."""
    gs = synth.shell_scene(1500, 2, seed=3, scale_mult=0.3)
    cams = {k_: v.to(cuda) for k_, v in synth.orbit_cameras(3, 80, 64, background=(0.1, 0.1, 0.1)).items()}
    gt_u8 = (synth.target_images(3, 64, 80, seed=2) * 255.0).round().to(torch.uint8)
    raw_a = {k_: v.to(cuda).requires_grad_() for k_, v in gs.items()}
    raw_b = {k_: v.to(cuda).requires_grad_() for k_, v in gs.items()}
    l1_a, mse_a = gs_utils.render_l1_loss(raw_a, cams, gt_u8.pin_memory())
    l1_b, mse_b = gs_utils.render_l1_loss(raw_b, cams, gt_u8.to(cuda).float() / 255.0)
    # the kernel divides by 255 (IEEE), torch multiplies by 1/255: last-bit differences in the targets
    assert torch.allclose(l1_a, l1_b, rtol=1e-6, atol=0) and torch.allclose(mse_a, mse_b, rtol=1e-6, atol=0)
    l1_a.sum().backward(); l1_b.sum().backward()
    for name in gs:
        H.assert_grad_close(raw_a[name].grad, raw_b[name].grad, name)


@pytest.mark.parametrize("n,size,views,scale_mult", [(4000, (160, 128), 3, 0.3), (900, (48, 40), 2, 1.0)])
def test_fused_tile_sort_is_bit_identical_to_oracle_tiling(cuda, sb_lib, n, size, views, scale_mult):
    """Counting sort by tile + per-tile (depth, id) sort == the oracle's stable 64-bit key sort:
    identical intersection count, gaussian_ids_sorted and tile_bins (north_star: bit-exact tiling)."""
    from splatformer_b200 import render
    gs = synth.shell_scene(n, 0, seed=17, scale_mult=scale_mult)
    gs["means"][: n // 50] = gs["means"][0]                    # exact depth ties: order must fall back to the id
    cams = synth.orbit_cameras(views, size[0], size[1])
    k = H.cam_scalars(cams)
    a = H.activated(gs)
    vms = torch.stack([R.viewmat_from_c2w(c)[:3] for c in cams["camera_to_worlds"]]).contiguous()
    intr = torch.tensor([k["fx"], k["fy"], k["cx"], k["cy"]]).repeat(views, 1)
    rgb, alpha = render.render_views(a["means"].to(cuda), a["scales"].to(cuda), a["quats"].to(cuda),
                                     a["opacities"].to(cuda).requires_grad_(), torch.sigmoid(gs["features_dc"]).to(cuda),
                                     -1, vms.to(cuda), None, intr.to(cuda), k["H"], k["W"], 16)
    ps = rgb.grad_fn.ps                                          # the pass's device state (render._Pass)
    gids_sorted, tile_bins = ps.gids_sorted.cpu(), ps.tile_bins.cpu()   # segment layout: tile t owns [t*cap, t*cap+count)
    tb = ((k["W"] + 15) // 16, (k["H"] + 15) // 16, 1)
    t = tb[0] * tb[1]
    total = 0
    for v in range(views):
        p = R.project_gaussians(a["means"], a["scales"], 1, a["quats"], vms[v], k["fx"], k["fy"], k["cx"], k["cy"],
                                k["H"], k["W"], 16)
        ni, cum = R.compute_cumulative_intersects(p[5])
        _, _, _, gs_ref, bins_ref = R.bin_and_sort_gaussians(n, ni, p[0], p[1], p[2], cum, tb, 16)
        mine = tile_bins[v * t:(v + 1) * t]
        assert torch.equal(mine[:, 1] - mine[:, 0], bins_ref[:, 1] - bins_ref[:, 0]), f"view {v}: tile counts differ"
        for tile in range(t):
            lo, hi = int(bins_ref[tile, 0]), int(bins_ref[tile, 1])
            if hi > lo:
                seg = gids_sorted[int(mine[tile, 0]):int(mine[tile, 1])]
                assert torch.equal(seg, gs_ref[lo:hi]), f"view {v} tile {tile}: sorted ids differ"
            else:
                assert int(mine[tile, 0]) == 0 and int(mine[tile, 1]) == 0
        total += ni
    assert render.last_num_intersects() == total


def test_fused_tile_overflow_falls_back_to_global_sort(cuda, sb_lib, monkeypatch):
    """A tile deeper than the segment capacity is detected on the device and the scene is re-binned
    with the global radix sort: same images, same gradients."""
    from splatformer_b200 import render
    gs = synth.shell_scene(3000, 1, seed=4, scale_mult=0.6)
    cams = {k_: v.to(cuda) for k_, v in synth.orbit_cameras(2, 96, 96).items()}
    gt = synth.target_images(2, 96, 96, seed=1).to(cuda)
    raw_a = {k_: v.to(cuda).requires_grad_() for k_, v in gs.items()}
    raw_b = {k_: v.to(cuda).requires_grad_() for k_, v in gs.items()}
    l1_a, _, img_a = gs_utils.render_l1_loss(raw_a, cams, gt, return_images=True)
    l1_a.sum().backward()
    tight = render.RenderPlan(tile_capacity=64)                  # this call site's plan: segments of 64 keys
    # first call of a shape through a plan: the flag is read right after the forward, the pass is repaired there
    with torch.no_grad():
        _, _, img_c = gs_utils.render_l1_loss(raw_b, cams, gt, return_images=True, plan=tight)
    assert tight.replays == 1 and torch.equal(img_a, img_c)
    assert tight.overflowed and tight.capacity(sb_lib, 2, 36, 3000) == 128      # the plan has learnt
    assert not render._default_plan.overflowed.get((2, 36, 3000)), "another call site's plan must not be touched"
    # later training steps are enqueued without a host synchronisation; 128 keys still overflow on this scene: the
    # backward notices, replays the forward with the global sort over the SAME image / loss-sum buffers, and goes on
    l1_b, _, img_b = gs_utils.render_l1_loss(raw_b, cams, gt, return_images=True, plan=tight)
    ps = img_b.grad_fn.ps
    l1_b.sum().backward()
    assert tight.replays == 2 and ps.rect_ids is None
    # the loss sums per rectangle on the rect-list path and per tile on the fallback: same terms, different order
    inv = 1.0 / (96 * 96 * 3)
    assert torch.equal(img_a, img_b) and torch.allclose(l1_a, ps.loss_sums[:, 0] * inv, rtol=2e-6, atol=0)
    for name in gs:
        H.assert_grad_close(raw_a[name].grad, raw_b[name].grad, name)


def test_per_view_wrapper_matches_fused_path(cuda, sb_lib):
    """gs_utils.per_view_multiimgs (the reference's loop on the three gsplat-compatible calls, what
    bench.py reports as wrapper_path) against rasterize_gaussians_to_multiimgs (fused)."""
    gs = synth.shell_scene(2000, 2, seed=14, scale_mult=0.3)
    cams = {k_: v.to(cuda) for k_, v in synth.orbit_cameras(3, 120, 88, background=(0.2, 0.3, 0.1)).items()}
    gt = synth.target_images(3, 88, 120, seed=9).to(cuda)
    raw_a = {k_: v.to(cuda).requires_grad_() for k_, v in gs.items()}
    raw_b = {k_: v.to(cuda).requires_grad_() for k_, v in gs.items()}
    rgbs_a, alphas_a = gs_utils.rasterize_gaussians_to_multiimgs(raw_a, cams)
    rgbs_b, alphas_b = gs_utils.per_view_multiimgs(raw_b, cams)
    for v in range(3):
        assert rgbs_b[v].shape == (88, 120, 3) and alphas_b[v].shape == (88, 120, 1)
        assert float((rgbs_a[v] - rgbs_b[v]).abs().max()) <= 1e-5
        assert float((alphas_a[v] - alphas_b[v]).abs().max()) <= 1e-5
    sum((r - g).abs().mean() for r, g in zip(rgbs_a, gt)).backward()
    sum((r - g).abs().mean() for r, g in zip(rgbs_b, gt)).backward()
    for name in gs:
        H.assert_grad_close(raw_a[name].grad, raw_b[name].grad, name)


def test_fused_scene_without_gaussians(cuda, sb_lib):
    """N = 0 (a scene whose Gaussians were all filtered out): background image, alpha 1 (the upstream
    empty-frame quirk), empty gradients, no error."""
    gs = synth.shell_scene(4, 1, seed=1)
    raw = {k_: v[:0].to(cuda).requires_grad_() for k_, v in gs.items()}
    cams = {k_: v.to(cuda) for k_, v in synth.orbit_cameras(2, 40, 24, background=(0.2, 0.4, 0.6)).items()}
    gt = torch.rand(2, 24, 40, 3, device=cuda)
    l1, mse, img = gs_utils.render_l1_loss(raw, cams, gt, return_images=True)
    assert torch.allclose(img, cams["background_color"].expand(2, 24, 40, 3))
    assert torch.allclose(l1, (img - gt).abs().mean(dim=(1, 2, 3)), rtol=1e-5)
    l1.sum().backward()
    assert raw["means"].grad is None or raw["means"].grad.shape == (0, 3)
    rgbs, alphas = gs_utils.rasterize_gaussians_to_multiimgs({k_: v.detach() for k_, v in raw.items()}, cams)
    assert bool((alphas[0] == 1).all()) and rgbs[1].shape == (24, 40, 3)


def test_fused_loss_under_grad_scaler_magnitudes(cuda, sb_lib):
    """AMP's GradScaler multiplies the upstream gradient by 2^16 (train.py:215,289): the kernels must not
    assume |v_out| <= 1 -- gradients scale exactly linearly, nothing overflows."""
    gs = synth.shell_scene(1500, 2, seed=6, scale_mult=0.3)
    cams = {k_: v.to(cuda) for k_, v in synth.orbit_cameras(2, 96, 64).items()}
    gt = synth.target_images(2, 64, 96, seed=3).to(cuda)
    raw_a = {k_: v.to(cuda).requires_grad_() for k_, v in gs.items()}
    raw_b = {k_: v.to(cuda).requires_grad_() for k_, v in gs.items()}
    l1_a, _ = gs_utils.render_l1_loss(raw_a, cams, gt)
    l1_b, _ = gs_utils.render_l1_loss(raw_b, cams, gt)
    l1_a.mean().backward()
    (l1_b.mean() * 65536.0).backward()
    for name in gs:
        assert torch.isfinite(raw_b[name].grad).all(), name
        assert H.rel_l2(raw_b[name].grad / 65536.0, raw_a[name].grad) <= 1e-5, name


def test_fused_backward_twice_and_without_grad(cuda, sb_lib):
    """The gradient records are zeroed during the forward (side stream); a second backward through a
    retained graph must start from zero again, and a forward under no_grad allocates none."""
    gs = synth.shell_scene(1200, 1, seed=8, scale_mult=0.3)
    cams = {k_: v.to(cuda) for k_, v in synth.orbit_cameras(2, 80, 64).items()}
    gt = synth.target_images(2, 64, 80, seed=5).to(cuda)
    raw = {k_: v.to(cuda).requires_grad_() for k_, v in gs.items()}
    l1, _ = gs_utils.render_l1_loss(raw, cams, gt)
    loss = l1.mean()
    loss.backward(retain_graph=True)
    first = {k_: p.grad.clone() for k_, p in raw.items()}
    for p in raw.values():
        p.grad = None
    loss.backward()
    for name in gs:
        assert torch.equal(first[name] != 0, raw[name].grad != 0), name
        assert H.rel_l2(raw[name].grad, first[name]) <= 1e-5, name      # float atomics: order may differ
    with torch.no_grad():
        l1_ng, _ = gs_utils.render_l1_loss(raw, cams, gt)
    assert torch.equal(l1_ng, l1.detach())



def test_fused_l1_plus_image_space_loss(cuda, sb_lib):
    """The reference trains with L1 + an image-space term (LPIPS, train.py:276-277): the images returned next to
    the fused L1 are differentiable, their gradient enters the raster backward beside sign(pred - gt).  Here the
    image-space term is 0.5 * mean((pred - gt)^2 * w): gradients against the float64 oracle."""
    gs = synth.shell_scene(1600, 2, seed=41, scale_mult=0.3)
    cams = synth.orbit_cameras(2, 88, 72, background=(0.05, 0.1, 0.2))
    gt = synth.target_images(2, 72, 88, seed=13)
    wgt = synth.target_images(2, 72, 88, seed=14)
    raw = {k_: v.to(cuda).requires_grad_() for k_, v in gs.items()}
    l1, mse, img = gs_utils.render_l1_loss(raw, {k_: v.to(cuda) for k_, v in cams.items()}, gt.to(cuda),
                                           return_images=True)
    assert img.requires_grad and not mse.requires_grad
    extra = 0.5 * ((img - gt.to(cuda)) ** 2 * wgt.to(cuda)).mean(dim=(1, 2, 3)).sum()
    (l1.sum() + extra).backward()
    ref, _ = H.oracle_f64_grads(gs, cams, gt, (0, 1),
                                extra=lambda rgb, v: 0.5 * ((rgb - gt[v].double()) ** 2 * wgt[v].double()).mean())
    for name in gs:
        assert H.rel_l2(raw[name].grad, ref[name]) <= H.GRAD_RTOL, (name, H.rel_l2(raw[name].grad, ref[name]))
    # the image-space term alone (no gradient on the fused L1 output)
    raw2 = {k_: v.to(cuda).requires_grad_() for k_, v in gs.items()}
    _, _, img2 = gs_utils.render_l1_loss(raw2, {k_: v.to(cuda) for k_, v in cams.items()}, gt.to(cuda), return_images=True)
    (0.5 * ((img2 - gt.to(cuda)) ** 2 * wgt.to(cuda)).mean(dim=(1, 2, 3)).sum()).backward()
    raw3 = {k_: v.to(cuda).requires_grad_() for k_, v in gs.items()}
    rgb3, _ = gs_utils.render_scene(raw3, **{k_: v.to(cuda) for k_, v in cams.items()})
    (0.5 * ((rgb3 - gt.to(cuda)) ** 2 * wgt.to(cuda)).mean(dim=(1, 2, 3)).sum()).backward()
    for name in gs:
        assert H.rel_l2(raw2[name].grad, raw3[name].grad) <= 5e-5, name      # (summation order of the atomics)


def test_render_views_argument_conventions(cuda, sb_lib):
    """opacities (N,) instead of (N,1): the gradient comes back in the caller's shape; world->camera matrices
    without camera centres: the SH view directions use the centres derived from the matrices."""
    from splatformer_b200 import render
    gs = synth.shell_scene(700, 2, seed=6, scale_mult=0.4)
    cams = synth.orbit_cameras(2, 64, 48)
    k = H.cam_scalars(cams)
    a = H.activated(gs, cuda)
    vms, campos = gs_utils.viewmats_from_c2w(cams["camera_to_worlds"].to(cuda))
    intr = torch.tensor([k["fx"], k["fy"], k["cx"], k["cy"]], device=cuda).repeat(2, 1)
    op2 = a["opacities"].clone().requires_grad_()
    op1 = a["opacities"].reshape(-1).clone().requires_grad_()
    rgb_a, _ = render.render_views(a["means"], a["scales"], a["quats"], op2, a["coeffs"], 2, vms, campos, intr, k["H"], k["W"])
    rgb_b, _ = render.render_views(a["means"], a["scales"], a["quats"], op1, a["coeffs"], 2, vms, None, intr, k["H"], k["W"])
    assert float((rgb_a - rgb_b).abs().max()) <= 1e-6          # derived centres: -R^T t vs the c2w translation
    rgb_a.sum().backward(); rgb_b.sum().backward()
    assert op1.grad.shape == (700,) and op2.grad.shape == (700, 1)
    assert H.rel_l2(op1.grad, op2.grad.reshape(-1)) <= 1e-5


def test_wrapper_edge_paths_on_the_fused_path(cuda, sb_lib):
    """gs_utils.py:47-51 / :72-76 through the kernels: a zero or NaN quaternion is rendered as (0,0,0,1) and gets
    no gradient, a mean on a camera centre is culled by the near plane; images and gradients equal the oracle's."""
    gs = synth.shell_scene(900, 1, seed=3, scale_mult=0.5)
    cams = synth.orbit_cameras(2, 64, 64)
    gs["quats"][5] = 0.0
    gs["quats"][6] = float("nan")
    gs["means"][7] = cams["camera_to_worlds"][0][:3, 3]
    gt = synth.target_images(2, 64, 64, seed=2)
    raw = {k_: v.to(cuda).requires_grad_() for k_, v in gs.items()}
    rgbs, alphas = gs_utils.rasterize_gaussians_to_multiimgs(raw, {k_: v.to(cuda) for k_, v in cams.items()})
    kw = {kk: cams[kk] for kk in ("cx", "cy", "fx", "fy", "width", "height", "background_color")}
    loss, masks = 0.0, []
    for v in range(2):
        rgb_ref, _, aux = R.render_view(gs, cams["camera_to_worlds"][v], return_aux=True, **kw)
        solid = aux["margin"] > H.FRAGILE_MARGIN
        assert torch.isfinite(rgbs[v]).all()
        assert float((rgbs[v].detach().cpu() - rgb_ref).abs().amax(-1)[solid].max()) <= H.IMG_ATOL
        masks.append(solid)
        loss = loss + ((rgbs[v] - gt[v].to(cuda)).abs() * solid[..., None].float().to(cuda)).mean()
    loss.backward()
    for name in gs:
        assert torch.isfinite(raw[name].grad).all(), name
    assert float(raw["quats"].grad[5].abs().sum()) == 0.0 and float(raw["quats"].grad[6].abs().sum()) == 0.0
    good = torch.ones(900, dtype=torch.bool); good[6] = False                  # (the NaN row itself is NaN in float64 too)
    ref, _ = H.oracle_f64_grads(gs, cams, gt, (0, 1), masks)
    for name in gs:
        H.assert_grad_close(raw[name].grad.cpu()[good], ref[name][good], name, rtol=H.GRAD_RTOL)
