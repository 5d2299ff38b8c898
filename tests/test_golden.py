"""Frozen golden vectors (tests/golden/*.npz, written by scripts/make_golden.py from the oracle).
CPU: the oracle still reproduces them (guards the oracle against drift).  GPU: the CUDA path,
through the drop-in wrapper and through the three gsplat-compatible calls, reproduces them --
tiling bit-exact, images within 1e-4, gradients within 1e-3 relative (north_star)."""
import os

import numpy as np
import pytest
import torch

from oracle import ref_torch as R
import helpers as H

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
NAMES = ["plumbing_4k_sh0_128", "small_sh3_ragged", "small_sh1_dense"]
GS_KEYS = ("means", "scales", "quats", "opacities", "features_dc", "features_rest")


def _load(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    gs = {k: torch.from_numpy(z[f"in_{k}"]) for k in GS_KEYS if f"in_{k}" in z}
    cams = {k[4:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("cam_")}
    return z, gs, cams


@pytest.mark.parametrize("name", NAMES)
def test_oracle_reproduces_golden(name):
    z, gs, cams = _load(name)
    views = cams["camera_to_worlds"].shape[0]
    h, w = int(cams["height"]), int(cams["width"])
    fx, fy, cx, cy = (float(cams[k]) for k in ("fx", "fy", "cx", "cy"))
    raw = {k: v.clone().requires_grad_() for k, v in gs.items()}
    gt = torch.from_numpy(z["gt"])
    loss = 0.0
    for v in range(views):
        rgb, alpha, aux = R.render_view(raw, cams["camera_to_worlds"][v], cx, cy, fx, fy, w, h,
                                        cams["background_color"], return_aux=True)
        assert aux["num_intersects"] == int(z[f"v{v}_num_intersects"])
        assert np.array_equal(aux["gaussian_ids_sorted"].numpy(), z[f"v{v}_gaussian_ids_sorted"])
        assert np.array_equal(aux["isect_ids_sorted"].numpy(), z[f"v{v}_isect_ids_sorted"])
        assert np.array_equal(aux["tile_bins"].numpy(), z[f"v{v}_tile_bins"])
        solid = torch.from_numpy(z[f"v{v}_solid"])
        assert float((rgb.detach() - torch.from_numpy(z[f"v{v}_rgb"])).abs()[solid].max()) <= 1e-5
        assert float((alpha.detach() - torch.from_numpy(z[f"v{v}_alpha"])).abs()[solid].max()) <= 1e-5
        loss = loss + ((rgb - gt[v]).abs() * solid[..., None]).mean()
    loss.backward()
    assert abs(float(loss.detach()) - float(z["loss"])) < 1e-6
    for k in gs:
        H.assert_grad_close(raw[k].grad, torch.from_numpy(z[f"grad_{k}"]), k, rtol=1e-4)


@pytest.mark.gpu
@pytest.mark.parametrize("name", NAMES)
def test_cuda_wrapper_reproduces_golden(cuda, sb_lib, name):
    from splatformer_b200 import gs_utils
    z, gs, cams = _load(name)
    views = cams["camera_to_worlds"].shape[0]
    raw = {k: v.to(cuda).requires_grad_() for k, v in gs.items()}
    rgbs, alphas = gs_utils.rasterize_gaussians_to_multiimgs(raw, {k: v.to(cuda) for k, v in cams.items()})
    gt = torch.from_numpy(z["gt"]).to(cuda)
    loss = 0.0
    for v in range(views):
        solid = torch.from_numpy(z[f"v{v}_solid"])
        err = (rgbs[v].detach().cpu() - torch.from_numpy(z[f"v{v}_rgb"])).abs().amax(-1)
        assert float(err[solid].max()) <= H.IMG_ATOL, (name, v, float(err[solid].max()))
        err_a = (alphas[v].detach().cpu() - torch.from_numpy(z[f"v{v}_alpha"])).abs()[..., 0]
        assert float(err_a[solid].max()) <= H.IMG_ATOL
        loss = loss + ((rgbs[v] - gt[v]).abs() * solid[..., None].to(cuda)).mean()
    loss.backward()
    assert abs(float(loss.detach()) - float(z["loss"])) < 1e-5
    for k in gs:          # against the float64 oracle's gradients of the same masked loss
        H.assert_grad_close(raw[k].grad, torch.from_numpy(z[f"grad64_{k}"]), k, rtol=H.GRAD_RTOL)


@pytest.mark.gpu
@pytest.mark.parametrize("name", NAMES)
def test_cuda_api_tiling_reproduces_golden_bit_exact(cuda, sb_lib, name):
    import gsplat
    z, gs, cams = _load(name)
    views = cams["camera_to_worlds"].shape[0]
    h, w = int(cams["height"]), int(cams["width"])
    fx, fy, cx, cy = (float(cams[k]) for k in ("fx", "fy", "cx", "cy"))
    n = gs["means"].shape[0]
    scales = torch.exp(gs["scales"])
    quats = gs["quats"] / gs["quats"].norm(dim=-1, keepdim=True)
    for v in range(views):
        vm = torch.from_numpy(z[f"v{v}_viewmat"]).to(cuda)
        got = gsplat.project_gaussians(gs["means"].to(cuda), scales.to(cuda), 1, quats.to(cuda), vm, fx, fy, cx, cy,
                                       h, w, 16)
        for nm, t in zip(("xys", "depths", "radii", "conics", "compensation", "num_tiles_hit", "cov3d"), got):
            assert np.array_equal(t.cpu().numpy(), z[f"v{v}_{nm}"]), f"{name} view {v}: {nm} not bit-identical"
        ni, cum = gsplat.compute_cumulative_intersects(got[5])
        assert ni == int(z[f"v{v}_num_intersects"])
        tb = ((w + 15) // 16, (h + 15) // 16, 1)
        _, _, keys_s, gids_s, bins = gsplat.bin_and_sort_gaussians(n, ni, got[0], got[1], got[2], cum, tb, 16)
        assert np.array_equal(keys_s.cpu().numpy(), z[f"v{v}_isect_ids_sorted"])
        assert np.array_equal(gids_s.cpu().numpy(), z[f"v{v}_gaussian_ids_sorted"])
        assert np.array_equal(bins.cpu().numpy(), z[f"v{v}_tile_bins"])
