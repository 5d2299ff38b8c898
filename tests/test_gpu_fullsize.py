"""BASELINE.json configs at their full sizes on the B200.

The oracle finishes one 100k-Gaussian 512x512 view in a few seconds, so config 2 is checked against
it directly on one view; everything larger goes through size-independent properties of the path:
  * the two binning strategies (fixed-capacity tile segments + per-tile sort + rect lists, and the
    global 64-bit radix sort + per-tile rasteriser) must give the same images and intersection count;
  * per-tile lists are sorted by (depth bits, id) and their lengths sum to sum(num_tiles_hit);
  * img(background b1) - img(background b0) = final_T * (b1 - b0)  (ties alpha to the colour path);
  * the backward is linear in the upstream gradient;
  * the fused path equals the per-view gsplat-compatible API.
"""
import pytest
import torch

import gsplat
from oracle import ref_torch as R
from splatformer_b200 import _abi, gs_utils, render, synth
import helpers as H

pytestmark = pytest.mark.gpu


def _dev(d, cuda, grad=False):
    return {k: (v.to(cuda).requires_grad_() if grad else v.to(cuda)) for k, v in d.items()}


def _force_global_sort(monkeypatch):
    monkeypatch.setattr(render, "DEFAULT_TILE_CAPACITY", 1 << 20)     # above the library's limit -> cap 0
    monkeypatch.setattr(render, "_default_plan", render.RenderPlan())


def _segments_sorted(gids_sorted, tile_bins, depths_v, v, tiles, n):
    """Every tile list of view v is ascending in (depth bits, id)."""
    bins = tile_bins[v * tiles:(v + 1) * tiles].long()
    cnt = bins[:, 1] - bins[:, 0]
    maxlen = int(cnt.max())
    if maxlen < 2:
        return True, int(cnt.sum())
    pos = torch.arange(maxlen, device=bins.device)
    idx = (bins[:, :1] + pos[None]).clamp_max(gids_sorted.numel() - 1)
    ids = gids_sorted[idx].long().clamp_(0, n - 1)
    key = (depths_v.view(torch.int32)[ids].long() << 32) | ids
    ok = (key[:, 1:] > key[:, :-1]) | (pos[None, 1:] >= cnt[:, None])
    return bool(ok.all()), int(cnt.sum())


def _oracle_projection(gs, cams, view, act_scales, act_quats):
    """The oracle's projection of one view from the activations the reference's wrapper would produce
    (torch.exp on the GPU, gs_utils.py:45; q / |q| with the (0,0,0,1) substitution, :46-51)."""
    k = H.cam_scalars(cams)
    vm = R.viewmat_from_c2w(cams["camera_to_worlds"][view])[:3].contiguous()
    with torch.no_grad():
        return R.project_gaussians(gs["means"], act_scales, 1, act_quats, vm, k["fx"], k["fy"], k["cx"], k["cy"],
                                   k["H"], k["W"], 16)


def _check_projection_bits(ps, view, proj, n, what):
    """depths / radii / tile counts / pixel centres of the fused pass against the oracle, bit for bit."""
    xys, depths, radii, conics, comp, tiles_hit, cov3d = proj
    assert torch.equal(ps.radii[view].cpu(), radii), f"{what} view {view}: radii differ"
    assert torch.equal(ps.tiles_hit[view].cpu(), tiles_hit), f"{what} view {view}: num_tiles_hit differ"
    drawn = radii > 0
    assert torch.equal(ps.depths[view].cpu()[drawn], depths[drawn]), f"{what} view {view}: depths differ"
    assert torch.equal(ps.packed[view, :, :2].cpu()[drawn], xys[drawn]), f"{what} view {view}: xys differ"
    return int(tiles_hit.sum()), int(drawn.sum())


def test_config2_benchmarked_call_against_oracle(cuda, sb_lib):
    """BASELINE config 2 through the very call bench.py times -- render_l1_loss on RAW parameters, OpenGL
    camera-to-world matrices and uint8 targets, all 32 views in one pass:
      * all 32 views: radii, tile counts, depths and pixel centres bit-identical to the oracle, the
        intersection count equal to the oracle's sum;
      * views 0, 10, 21, 31: the sorted Gaussian ids of every tile identical to the oracle's, image and
        transmittance within 1e-4 (pixels whose threshold decisions are closer than 2e-5 to a threshold are
        set aside and counted), per-view loss within 1e-5;
      * gradients of the four views' losses w.r.t. the raw parameters within 1e-3 of the FLOAT64 oracle."""
    name = "objaverse_100k_sh3_512"
    gs, cams, gt = synth.make_config(name, 0)
    V, n = cams["camera_to_worlds"].shape[0], gs["means"].shape[0]
    gt_u8 = (gt * 255.0).round().to(torch.uint8)
    raw = _dev(gs, cuda, grad=True)
    l1, mse, img = gs_utils.render_l1_loss(raw, _dev(cams, cuda), gt_u8.to(cuda), return_images=True)
    ps = img.grad_fn.ps
    T = ps.num_tiles
    act_scales = torch.exp(gs["scales"].cuda()).cpu()
    act_quats = R.wrapper_normalise_quats(gs["quats"])
    total = 0
    for v in range(V):
        cnt, _ = _check_projection_bits(ps, v, _oracle_projection(gs, cams, v, act_scales, act_quats), n, name)
        bins = ps.tile_bins[v * T:(v + 1) * T]
        assert int((bins[:, 1] - bins[:, 0]).sum()) == cnt, f"view {v}: tile lists hold a different number of entries"
        total += cnt
    assert render.last_num_intersects() == total
    H.report(config=name, check="projection bits + intersection count, raw-parameter fused call", views=V,
             pairs=V * n, intersections=total, mismatches=0)

    sel = (0, 10, 21, 31)
    kw = {k: cams[k] for k in ("cx", "cy", "fx", "fy", "width", "height", "background_color")}
    gt_f = gt_u8.float() / 255.0
    raw64 = {k: v.double().requires_grad_() for k, v in gs.items()}
    loss64 = 0.0
    for v in sel:
        with torch.no_grad():
            rgb_ref, a_ref, aux = R.render_view(gs, cams["camera_to_worlds"][v], return_aux=True, exp=H.gpu_exp,
                                                sigmoid=H.gpu_sigmoid, **kw)
        mine, cnt = H.tile_lists_from_segments(ps.gids_sorted, ps.tile_bins, v, T)
        ref_bins = aux["tile_bins"].long()
        assert torch.equal(cnt, ref_bins[:, 1] - ref_bins[:, 0]), f"view {v}: per-tile counts differ"
        assert torch.equal(mine, aux["gaussian_ids_sorted"]), f"view {v}: sorted Gaussian ids differ"
        assert aux["num_intersects"] == int(cnt.sum())
        solid = aux["margin"] > H.FRAGILE_MARGIN
        frac = 1.0 - float(solid.float().mean())
        err = (img[v].detach().cpu() - rgb_ref).abs().amax(-1)
        err_t = ((1.0 - ps.final_ts[v].cpu()) - a_ref[..., 0]).abs()
        H.report(config=name, check="image / alpha vs f32 oracle", view=v, pixels=solid.numel(),
                 set_aside_fraction=frac, max_err_solid=float(err[solid].max()), max_err_all=float(err.max()),
                 sorted_ids=int(cnt.sum()))
        assert frac <= 2e-3
        assert float(err[solid].max()) <= H.IMG_ATOL and float(err_t[solid].max()) <= H.IMG_ATOL
        assert abs(float(l1[v]) - float((rgb_ref - gt_f[v]).abs().mean())) <= 1e-5
        rgb64, _ = R.render_view(raw64, cams["camera_to_worlds"][v].double(), **kw)
        loss64 = loss64 + (rgb64 - gt_f[v].double()).abs().mean()
    w = torch.zeros(V, device=cuda)
    w[list(sel)] = 1.0
    (l1 * w).sum().backward()
    loss64.backward()
    # the float32 ORACLE against the float64 oracle, same metric: what binary32 arithmetic itself costs here
    raw32 = {k: v.clone().requires_grad_() for k, v in gs.items()}
    loss32 = 0.0
    for v in sel:
        rgb32, _ = R.render_view(raw32, cams["camera_to_worlds"][v], exp=H.gpu_exp, sigmoid=H.gpu_sigmoid, **kw)
        loss32 = loss32 + (rgb32 - gt_f[v]).abs().mean()
    loss32.backward()
    for key in gs:
        got, ref, o32 = raw[key].grad.double().cpu(), raw64[key].grad, raw32[key].grad.double()
        rms = float(ref.pow(2).mean().sqrt())
        band = H.GRAD_RTOL * (ref.abs() + rms)
        out_k, out_o = (got - ref).abs() > band, (o32 - ref).abs() > band
        H.report(config=name, check="gradient vs f64 oracle (fused L1 of 4 views, raw parameters)", tensor=key,
                 rel_l2_kernel=H.rel_l2(got, ref), rel_l2_f32_oracle=H.rel_l2(o32, ref), elements=ref.numel(),
                 outside_band_kernel=int(out_k.sum()), outside_band_f32_oracle=int(out_o.sum()),
                 worst_kernel=float(((got - ref).abs() / (ref.abs() + rms)).max()),
                 worst_f32_oracle=float(((o32 - ref).abs() / (ref.abs() + rms)).max()))
        # north_star: gradients within 1e-3 relative.  Relative L2 per tensor; elementwise, the few entries outside
        # the 1e-3 band are Gaussians fed by a pixel whose alpha >= 1/255 / T <= 1e-4 decision flips between
        # binary32 and binary64 (the float32 oracle shows the same number of them): counted, bounded, reported
        assert H.rel_l2(got, ref) <= H.GRAD_RTOL, key
        assert int(out_k.sum()) <= max(2 * int(out_o.sum()), int(1e-3 * ref.numel())), key


def _tile_subset_check(name, gs, cams, views, ps, img, alpha, tiles_per_view, seed, grad=None):
    """Shared by configs 4 and 5: full-N projection bits, then `tiles_per_view` random occupied tiles per
    view -- sorted ids identical to the oracle's, image and alpha within 1e-4 on those tiles' pixels.
    Returns the oracle loss over the chosen tiles (mean |rgb - gt| on their pixels) if `grad` = (raw_ref, gt)."""
    n = gs["means"].shape[0]
    k = H.cam_scalars(cams)
    T = ps.num_tiles
    tb = (ps.tiles_x, ps.tiles_y, 1)
    act_scales = torch.exp(gs["scales"].cuda()).cpu()
    act_quats = R.wrapper_normalise_quats(gs["quats"])
    act_opac = torch.sigmoid(gs["opacities"].cuda()).cpu()
    gen = torch.Generator().manual_seed(seed)
    total, masks, loss_ref = 0, [], 0.0
    for v in views:
        proj = _oracle_projection(gs, cams, v, act_scales, act_quats)
        cnt, drawn = _check_projection_bits(ps, v, proj, n, name)
        total += cnt
        bins = ps.tile_bins[v * T:(v + 1) * T].long().cpu()
        assert int((bins[:, 1] - bins[:, 0]).sum()) == cnt
        occupied = torch.nonzero(bins[:, 1] > bins[:, 0]).flatten()
        pick = occupied[torch.randperm(occupied.numel(), generator=gen)[:tiles_per_view]].tolist()
        xys, depths, radii, conics = proj[0], proj[1], proj[2], proj[3]
        gids, tbins = R.tile_lists(xys, depths, radii, tb, 16, pick)
        entries = 0
        gsorted = ps.gids_sorted.cpu() if ps.gids_sorted.numel() < (1 << 28) else None
        for t in pick:
            lo, hi = int(bins[t, 0]), int(bins[t, 1])
            seg = gsorted[lo:hi] if gsorted is not None else ps.gids_sorted[lo:hi].cpu()
            assert torch.equal(seg, gids[int(tbins[t, 0]):int(tbins[t, 1])]), f"{name} view {v} tile {t}: sorted ids differ"
            entries += hi - lo
        # colours and compositing of the chosen tiles
        c2w = cams["camera_to_worlds"][v]
        if grad is not None:
            raw_ref, gt = grad
            kw = {kk: cams[kk] for kk in ("cx", "cy", "fx", "fy", "width", "height", "background_color")}
            rgb_ref, a_ref, aux = R.render_view(raw_ref, c2w, return_aux=True, exp=H.gpu_exp, sigmoid=H.gpu_sigmoid,
                                                tiles=pick, **kw)
            margin = aux["margin"]
        else:
            coeffs = torch.cat([gs["features_dc"].unsqueeze(1), gs["features_rest"]], dim=1)
            dirs = gs["means"] - c2w[:3, 3]
            dirs = dirs / dirs.norm(dim=-1, keepdim=True)
            cols = torch.clamp(R.spherical_harmonics(3, dirs, coeffs) + 0.5, min=0.0)
            out, t_img, _, margin = R.rasterize_forward(xys, conics, cols, act_opac, gids, tbins, k["H"], k["W"], 16,
                                                        cams["background_color"], tiles=pick)
            rgb_ref, a_ref = out.clamp(max=1.0), (1.0 - t_img)[..., None]
        mask = torch.zeros(k["H"], k["W"], dtype=torch.bool)
        for t in pick:
            ty, tx = divmod(t, tb[0])
            mask[ty * 16:ty * 16 + 16, tx * 16:tx * 16 + 16] = True
        solid = mask & (margin > H.FRAGILE_MARGIN)
        frac = 1.0 - float(solid.sum()) / float(mask.sum())
        err = (img[v].detach().cpu() - rgb_ref.detach()).abs().amax(-1)
        err_a = (alpha[v].detach().cpu().reshape(k["H"], k["W"]) - a_ref.detach()[..., 0]).abs()
        H.report(config=name, check="tile subset: sorted ids + image / alpha vs oracle", view=v, tiles=len(pick),
                 list_entries=entries, pixels=int(mask.sum()), set_aside_fraction=frac,
                 max_err_solid=float(err[solid].max()), max_err_all=float(err[mask].max()),
                 gaussians_drawn=drawn, intersections=cnt)
        # (lists here are thousands deep: two orders of magnitude more threshold decisions per pixel than config 2)
        assert frac <= 2e-2
        assert float(err[solid].max()) <= H.IMG_ATOL and float(err_a[solid].max()) <= H.IMG_ATOL
        masks.append(mask)
        if grad is not None:
            loss_ref = loss_ref + ((rgb_ref - gt[v]).abs() * mask[..., None]).sum() / (3.0 * float(mask.sum()))
    return total, masks, loss_ref


def test_config2_full_scene_binning_paths_agree(cuda, sb_lib, monkeypatch):
    """All 32 views: tile-segment sort + rect lists == global radix sort + per-tile rasteriser."""
    gs, cams, gt = synth.make_config("objaverse_100k_sh3_512", 1)
    cams_d, gt_d = _dev(cams, cuda), gt.to(cuda)
    w = torch.linspace(0.5, 1.5, gt.shape[0], device=cuda)
    raw_a, raw_b, raw_c = _dev(gs, cuda, True), _dev(gs, cuda, True), _dev(gs, cuda, True)
    l1_a, mse_a, img_a = gs_utils.render_l1_loss(raw_a, cams_d, gt_d, return_images=True)
    isect_a = render.last_num_intersects()
    (l1_a * w).sum().backward()
    # linearity of the backward in the upstream gradient
    l1_c, _ = gs_utils.render_l1_loss(raw_c, cams_d, gt_d)
    (l1_c * (2.0 * w)).sum().backward()
    for name in gs:
        assert H.rel_l2(raw_c[name].grad, 2.0 * raw_a[name].grad) <= 1e-5, name
    _force_global_sort(monkeypatch)
    l1_b, mse_b, img_b = gs_utils.render_l1_loss(raw_b, cams_d, gt_d, return_images=True)
    assert render.last_num_intersects() == isect_a and isect_a > 32 * 100_000
    assert torch.equal(img_a, img_b)
    assert torch.allclose(l1_a, l1_b, rtol=1e-6, atol=0) and torch.allclose(mse_a, mse_b, rtol=1e-6, atol=0)
    (l1_b * w).sum().backward()
    for name in gs:
        assert torch.isfinite(raw_a[name].grad).all()
        H.assert_grad_close(raw_a[name].grad, raw_b[name].grad, name)


def test_config2_tile_lists_sorted_and_counted(cuda, sb_lib):
    gs, cams, _ = synth.make_config("objaverse_100k_sh3_512", 2)
    a = H.activated(gs, cuda)
    k = H.cam_scalars(cams)
    V, n = cams["camera_to_worlds"].shape[0], a["means"].shape[0]
    vms, campos = gs_utils.viewmats_from_c2w(cams["camera_to_worlds"].to(cuda))
    intr = torch.tensor([k["fx"], k["fy"], k["cx"], k["cy"]], device=cuda).repeat(V, 1)
    bg0, bg1 = torch.zeros(3, device=cuda), torch.tensor([1.0, 0.5, 0.25], device=cuda)
    args = (a["means"], a["scales"], a["quats"], a["opacities"].requires_grad_(), a["coeffs"], 3, vms, campos, intr,
            k["H"], k["W"], 16)
    rgb0, alpha0 = render.render_views(*args, background=bg0)
    ps = rgb0.grad_fn.ps
    gids_sorted, tile_bins = ps.gids_sorted, ps.tile_bins
    tiles = ((k["W"] + 15) // 16) * ((k["H"] + 15) // 16)
    total = 0
    for v in (0, 7, 31):
        p = gsplat.project_gaussians(a["means"], a["scales"], 1, a["quats"], vms[v], k["fx"], k["fy"], k["cx"],
                                     k["cy"], k["H"], k["W"], 16)
        ok, cnt = _segments_sorted(gids_sorted, tile_bins, p[1].contiguous(), v, tiles, n)
        assert ok, f"view {v}: a tile list is not ascending in (depth, id)"
        assert cnt == int(p[5].sum()), f"view {v}: list lengths do not sum to sum(num_tiles_hit)"
        total += cnt
    assert total > 0
    # background identity: the colour path and alpha agree on the transmittance
    rgb1, alpha1 = render.render_views(*args, background=bg1)
    assert torch.equal(alpha0, alpha1)
    assert float(alpha0.min()) >= 0.0 and float(alpha0.max()) <= 1.0
    t_from_rgb = (rgb1 - rgb0).detach() / bg1
    assert float((t_from_rgb - (1.0 - alpha0.detach())[..., None]).abs().max()) <= 2e-6


def test_config3_scenes_are_independent_units(cuda, sb_lib):
    """Train-step shape (several scenes per step): a scene's loss and gradients do not depend on what
    was rendered before it (no state leaks between scenes through the library or the host wrapper)."""
    outs = []
    for order in ((0, 1, 2), (2, 0, 1)):
        res = {}
        for s in order:
            gs, cams, gt = synth.make_config("objaverse_100k_sh3_512", s, num_views=8)
            raw = _dev(gs, cuda, True)
            l1, _ = gs_utils.render_l1_loss(raw, _dev(cams, cuda), gt.to(cuda))
            l1.mean().backward()
            res[s] = (l1.detach().clone(), raw["means"].grad.clone(), raw["features_rest"].grad.clone())
        outs.append(res)
    for s in (0, 1, 2):
        assert torch.equal(outs[0][s][0], outs[1][s][0])
        assert H.rel_l2(outs[0][s][1], outs[1][s][1]) <= 1e-5          # atomics: summation order varies
        assert H.rel_l2(outs[0][s][2], outs[1][s][2]) <= 1e-5


def test_config4_forward_1m_gaussians_1080p(cuda, sb_lib, monkeypatch):
    """Real-OOD-shaped eval scene, forward only (no autograd graph): full-N projection bits and 64 random
    tiles per view (sorted ids, image, alpha) against the oracle; then the other checks of the shape."""
    name = "real_1m_sh3_1080p"
    gs, cams, _ = synth.make_config(name, 0, num_views=2)
    raw, cams_d = _dev(gs, cuda), _dev(cams, cuda)
    monkeypatch.setattr(render, "KEEP_LAST_PASS", True)
    with torch.no_grad():
        rgbs, alphas = gs_utils.rasterize_gaussians_to_multiimgs(raw, cams_d)
        ps = render._last_pass
        isect = render.last_num_intersects()
        assert rgbs[0].shape == (1080, 1920, 3) and alphas[0].shape == (1080, 1920, 1)
        assert isect > 1_000_000
        total, _, _ = _tile_subset_check(name, gs, cams, (0, 1), ps, rgbs, alphas, 64, seed=40)
        assert total == isect
        for im, al in zip(rgbs, alphas):
            assert torch.isfinite(im).all() and float(im.max()) <= 1.0 and float(im.min()) >= 0.0
            assert float(al.min()) >= 0.0 and float(al.max()) <= 1.0 and float(al.mean()) > 0.2
        # per-view gsplat-compatible API on view 1
        k = H.cam_scalars(cams)
        c2w = cams_d["camera_to_worlds"][1]
        vm = gs_utils.viewmats_from_c2w(c2w[None])[0][0]
        a = H.activated(gs, cuda)
        dirs = a["means"] - c2w[:3, 3]
        dirs = dirs / dirs.norm(dim=-1, keepdim=True)
        cols = torch.clamp(gsplat.spherical_harmonics(3, dirs, a["coeffs"]) + 0.5, min=0.0)
        xys, depths, radii, conics, comp, tiles, cov3d = gsplat.project_gaussians(
            a["means"], a["scales"], 1, a["quats"], vm, k["fx"], k["fy"], k["cx"], k["cy"], k["H"], k["W"], 16)
        rgb1, alpha1 = gsplat.rasterize_gaussians(xys, depths, radii, conics, tiles, cols, a["opacities"], k["H"],
                                                  k["W"], 16, background=cams_d["background_color"],
                                                  return_alpha=True)
        # the per-view API gets torch's quaternion norm and camera inversion (last-bit differences in the
        # projected Gaussians), so among 2M pixels a few alpha >= 1/255 decisions flip: each flip moves
        # a pixel by at most ~1/255 of a colour difference; everything else agrees to rounding
        for err in ((torch.clamp(rgb1, max=1.0) - rgbs[1]).abs().amax(-1), (alpha1 - alphas[1][..., 0]).abs()):
            assert float((err > 2e-5).float().mean()) <= 1e-4
            assert float(err.max()) <= 1e-2
        # the other binning strategy
        _force_global_sort(monkeypatch)
        rgbs_b, alphas_b = gs_utils.rasterize_gaussians_to_multiimgs(raw, cams_d)
        assert render.last_num_intersects() == isect
        for x, y in zip(rgbs + alphas, rgbs_b + alphas_b):
            assert torch.equal(x, y)
    monkeypatch.setattr(render, "_last_pass", None)


def test_config5_stress_6m_gaussians_4k(cuda, sb_lib):
    """6M Gaussians with ~60 px screen radii at 3840x2160 (sort- and intersection-dominated), forward +
    backward, TWO views.  Tiles are far deeper than any segment capacity: the
    device-side overflow flag must route the scene to the global radix sort.  Against the oracle: full-N
    projection bits, 48 random tiles per view (sorted ids, image, alpha), and the gradients of the L1 loss
    restricted to those tiles' pixels w.r.t. all 6M Gaussians' raw parameters."""
    name = "stress_6m_4k"
    gs, cams, gt = synth.make_config(name, 0, num_views=2)
    raw, cams_d = _dev(gs, cuda, True), _dev(cams, cuda)
    gt_u8 = (gt * 255.0).round().to(torch.uint8).to(cuda)
    plan = render.RenderPlan()
    l1, mse, img = gs_utils.render_l1_loss(raw, cams_d, gt_u8, return_images=True, plan=plan)
    # first call of the shape through this plan: the overflow flag is read right after the forward and the pass
    # is redone with the global sort before anything is returned
    assert plan.replays == 1
    l1.sum().backward()
    isect = render.last_num_intersects()
    assert isect > 100_000_000
    assert torch.isfinite(img).all() and float(img.min()) >= 0.0 and float(img.max()) <= 1.0
    ref_l1 = (img.detach() - gt_u8.float() / 255.0).abs().mean(dim=(1, 2, 3))
    assert torch.allclose(l1.detach(), ref_l1, rtol=1e-5, atol=1e-7)
    g1 = {k: v.grad.clone() for k, v in raw.items()}
    for key, g in g1.items():
        assert torch.isfinite(g).all(), key
    assert float(g1["means"].abs().sum()) > 0 and float(g1["features_rest"].abs().sum()) > 0
    # second pass (the plan now knows the shape: straight to the global sort): linearity of the backward
    raw2 = _dev(gs, cuda, True)
    l1b, _, imgb = gs_utils.render_l1_loss(raw2, cams_d, gt_u8, return_images=True, plan=plan)
    assert plan.replays == 1 and render.last_num_intersects() == isect and torch.equal(img, imgb)
    (2.0 * l1b).sum().backward()
    for key in gs:
        assert H.rel_l2(raw2[key].grad, 2.0 * g1[key]) <= 1e-4, key
    del raw2, l1b, imgb
    # third pass: an image-space loss on a random subset of tiles (through the differentiable images of the
    # same fused call) against the oracle compositing exactly those tiles
    raw3 = _dev(gs, cuda, True)
    _, _, img3 = gs_utils.render_l1_loss(raw3, cams_d, gt_u8, return_images=True, plan=plan)
    ps = img3.grad_fn.ps
    raw_ref = {k: v.clone().requires_grad_() for k, v in gs.items()}
    gt_f = (gt_u8.float() / 255.0).cpu()
    alpha = [(1.0 - ps.final_ts[v]) for v in range(2)]
    total, masks, loss_ref = _tile_subset_check(name, gs, cams, (0, 1), ps, img3, alpha, 48, seed=50,
                                                grad=(raw_ref, gt_f))
    assert total == isect
    m = torch.stack(masks).to(cuda)[..., None].float()
    loss = (((img3 - gt_u8.float() / 255.0).abs() * m).sum(dim=(1, 2, 3)) / (3.0 * m.sum(dim=(1, 2, 3)))).sum()
    assert abs(float(loss) - float(loss_ref)) <= 2e-5
    loss.backward()
    loss_ref.backward()
    for key in gs:
        got, ref = raw3[key].grad, raw_ref[key].grad
        H.report(config=name, check="gradient of the tile-subset L1 loss vs f32 oracle (6M Gaussians)", tensor=key,
                 rel_l2=H.rel_l2(got, ref), norm_ref=float(ref.norm()))
        assert H.rel_l2(got, ref) <= H.GRAD_RTOL, key
    del raw3
    del raw, g1
    torch.cuda.empty_cache()


def test_stress_shaped_small_case_against_oracle(cuda, sb_lib):
    """Same regime as config 5 (every Gaussian covers tens of tiles, tiles hundreds deep, early
    termination everywhere) at a size the oracle handles."""
    n = 2500
    gs = synth.shell_scene(n, 1, seed=77, scale_sigma=0.25)
    gs["scales"] = gs["scales"] * 0 + torch.log(torch.tensor(0.02)) + 0.25 * torch.randn(n, 3, generator=torch.Generator().manual_seed(5))
    cams = synth.orbit_cameras(2, 160, 96, background=(0.0, 0.0, 0.0))
    gt = synth.target_images(2, 96, 160, seed=6)
    raw = _dev(gs, cuda, True)
    raw_ref = {k: v.clone().requires_grad_() for k, v in gs.items()}
    rgbs, alphas = gs_utils.rasterize_gaussians_to_multiimgs(raw, _dev(cams, cuda))
    assert render.last_num_intersects() > 10 * n
    loss, masks = 0.0, []
    for v in range(2):
        rgb_ref, a_ref, aux = R.render_view(raw_ref, cams["camera_to_worlds"][v], cams["cx"], cams["cy"], cams["fx"],
                                            cams["fy"], cams["width"], cams["height"], cams["background_color"],
                                            return_aux=True)
        solid = aux["margin"] > H.FRAGILE_MARGIN
        assert float((rgbs[v].detach().cpu() - rgb_ref.detach()).abs().amax(-1)[solid].max()) <= H.IMG_ATOL
        assert float((alphas[v].detach().cpu() - a_ref.detach()).abs()[solid].max()) <= H.IMG_ATOL
        m = solid[..., None].float()
        masks.append(solid)
        loss = loss + ((rgbs[v] - gt[v].to(cuda)).abs() * m.to(cuda)).mean()
    loss.backward()
    ref, _ = H.oracle_f64_grads(gs, cams, gt, (0, 1), masks)
    for name in gs:
        H.assert_grad_close(raw[name].grad, ref[name], name, rtol=H.GRAD_RTOL)
