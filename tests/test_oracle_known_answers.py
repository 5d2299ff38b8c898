"""Known-answer pins for the oracle (oracle/ref_torch.py).  The reference owns no golden vectors for
the gsplat boundary, so these analytic cases are what fixes the oracle's
constants: SH basis values, the EWA projection of an isotropic Gaussian, the 0.3 px blur, the
1/255 and 1e-4 thresholds, depth ordering, the 0.999 clamp and the empty-frame quirk."""
import math

import pytest
import torch

from oracle import ref_torch as R
from splatformer_b200 import synth
import helpers as H


def _identity_view():
    vm = torch.eye(4)[:3].contiguous()      # camera at origin looking down +z (OpenCV)
    return vm


def _one_gaussian(z=2.0, s=0.05, xy=(0.0, 0.0)):
    means = torch.tensor([[xy[0], xy[1], z]])
    scales = torch.full((1, 3), s)
    quats = torch.tensor([[1.0, 0.0, 0.0, 0.0]])
    return means, scales, quats


def test_sh_constants_and_axis_values():
    d = torch.tensor([[0.0, 0.0, 1.0], [1.0, 0.0, 0.0], [0.0, 1.0, 0.0], [0.0, 0.0, -2.0]])
    Y = R.sh_basis(4, d)
    assert Y.shape == (4, 25)
    assert torch.allclose(Y[:, 0], torch.full((4,), 0.5 * math.sqrt(1.0 / math.pi)))
    c1 = math.sqrt(3.0 / (4.0 * math.pi))
    # (x,y,z) = +z: Y1 = -c1*y = 0, Y2 = c1*z = c1, Y3 = -c1*x = 0
    assert torch.allclose(Y[0, 1:4], torch.tensor([0.0, c1, 0.0]))
    assert torch.allclose(Y[1, 1:4], torch.tensor([0.0, 0.0, -c1]))
    assert torch.allclose(Y[2, 1:4], torch.tensor([-c1, 0.0, 0.0]))
    # the direction is normalised inside: (0,0,-2) == (0,0,-1)
    assert torch.allclose(Y[3, 1:4], torch.tensor([0.0, -c1, 0.0]))
    # Y6 = 1/4 sqrt(5/pi) (3z^2 - 1) at +z
    assert abs(float(Y[0, 6]) - 0.25 * math.sqrt(5.0 / math.pi) * 2.0) < 1e-6
    # Y20 (l=4, m=0) at +z = 3/16 sqrt(1/pi) * (35 - 30 + 3)
    assert abs(float(Y[0, 20]) - (3.0 / 16.0) * math.sqrt(1.0 / math.pi) * 8.0) < 1e-6


def test_sh_orthonormal_on_sphere():
    # Monte-Carlo-free check: Lebedev-like product grid, integral Y_i Y_j dOmega = delta_ij
    nt, nphi = 64, 128
    x, w = (torch.from_numpy(a) for a in __import__("numpy").polynomial.legendre.leggauss(nt))
    phi = (torch.arange(nphi, dtype=torch.float64) + 0.5) * (2.0 * math.pi / nphi)
    ct = x[:, None].expand(nt, nphi)
    st = torch.sqrt(1.0 - ct * ct)
    d = torch.stack([st * torch.cos(phi)[None, :], st * torch.sin(phi)[None, :], ct], dim=-1).reshape(-1, 3)
    wt = (w[:, None] * (2.0 * math.pi / nphi)).expand(nt, nphi).reshape(-1)
    Y = R.sh_basis(4, d)
    gram = (Y * wt[:, None]).T @ Y
    assert torch.allclose(gram, torch.eye(25, dtype=torch.float64), atol=1e-10)


def test_sh_degrees_to_use_and_grad():
    torch.manual_seed(0)
    d = torch.randn(10, 3)
    c = torch.randn(10, 16, 3, requires_grad=True)
    out = R.spherical_harmonics(1, d, c)
    out.sum().backward()
    assert bool((c.grad[:, 4:, :] == 0).all()) and bool((c.grad[:, :4, :] != 0).any())
    Y = R.sh_basis(1, d)
    assert torch.allclose(c.grad[:, :4, 0], Y)
    with pytest.raises(AssertionError):
        R.spherical_harmonics(2, d, torch.randn(10, 4, 3))


def test_projection_isotropic_gaussian_centre():
    fx = fy = 100.0
    W = H = 64
    z, s = 2.0, 0.05
    means, scales, quats = _one_gaussian(z, s)
    xys, depths, radii, conics, comp, tiles, cov3d = R.project_gaussians(
        means, scales, 1, quats, _identity_view(), fx, fy, 32.0, 32.0, H, W, 16)
    var = (fx * s / z) ** 2                       # EWA: (f s / z)^2, then + 0.3 blur
    assert torch.allclose(xys, torch.tensor([[32.0, 32.0]]), atol=1e-4)
    assert float(depths) == z
    assert torch.allclose(cov3d, torch.tensor([[s * s, 0, 0, s * s, 0, s * s]]), atol=1e-9)
    assert torch.allclose(conics, torch.tensor([[1.0 / (var + 0.3), 0.0, 1.0 / (var + 0.3)]]), rtol=1e-5)
    assert abs(float(comp) - var / (var + 0.3)) < 1e-5
    # radius = ceil(3 sqrt(lambda_max)); lambda = b + sqrt(max(0.1, b^2 - det)) with b^2 - det = 0 -> 0.1
    lam = (var + 0.3) + math.sqrt(0.1)
    assert int(radii) == math.ceil(3.0 * math.sqrt(lam))
    # centre (32,32) with radius r: tiles [ (32-r)/16 , (32+r)/16 + 1 )
    r = int(radii)
    lo, hi = int((32 - r) / 16), min(int((32 + r) / 16 + 1), 4)
    assert int(tiles) == (hi - lo) ** 2


def test_projection_culls():
    fx = fy = 100.0
    vm = _identity_view()
    m, s, q = _one_gaussian(z=0.005)             # behind the near plane (clip 0.01)
    out = R.project_gaussians(m, s, 1, q, vm, fx, fy, 32.0, 32.0, 64, 64, 16)
    assert int(out[2]) == 0 and int(out[5]) == 0 and bool((out[6] == 0).all())
    m, s, q = _one_gaussian(z=2.0, xy=(50.0, 0.0))   # far outside the image: no tile, but cov3d written
    out = R.project_gaussians(m, s, 1, q, vm, fx, fy, 32.0, 32.0, 64, 64, 16)
    assert int(out[2]) == 0 and int(out[5]) == 0 and float(out[0].abs().sum()) == 0.0
    assert float(out[6][0, 0]) > 0
    with pytest.raises(AssertionError):
        R.project_gaussians(m, s, 1, q, vm, fx, fy, 32.0, 32.0, 64, 64, 17)


def _raster_single(opac, color, bg, z=2.0, s=0.05, size=32):
    fx = fy = 100.0
    c = size / 2.0
    m, sc, q = _one_gaussian(z, s)
    xys, depths, radii, conics, comp, tiles, _ = R.project_gaussians(m, sc, 1, q, _identity_view(), fx, fy, c, c,
                                                                     size, size, 16)
    img, alpha = R.rasterize_gaussians(xys, depths, radii, conics, tiles, torch.tensor([color]),
                                       torch.tensor([[opac]]), size, size, 16, background=torch.tensor(bg),
                                       return_alpha=True)
    return img, alpha, conics, radii


def test_raster_single_gaussian_analytic():
    opac, color, bg = 0.8, [0.9, 0.5, 0.1], [0.2, 0.3, 0.4]
    size = 32
    img, alpha, conics, radii = _raster_single(opac, color, bg, size=size)
    var = (100.0 * 0.05 / 2.0) ** 2 + 0.3
    ys, xs = torch.meshgrid(torch.arange(size) + 0.5, torch.arange(size) + 0.5, indexing="ij")
    d2 = (xs - 16.0) ** 2 + (ys - 16.0) ** 2
    a = opac * torch.exp(-0.5 * d2 / var)
    a = torch.where(a >= 1.0 / 255.0, a, torch.zeros_like(a))
    # tiles outside the Gaussian's bounding box see nothing (radius-limited support)
    r = int(radii)
    tlo, thi = int((16 - r) / 16), min(int((16 + r) / 16 + 1), 2)
    inside = (xs >= tlo * 16) & (xs < thi * 16) & (ys >= tlo * 16) & (ys < thi * 16)
    a = torch.where(inside, a, torch.zeros_like(a))
    assert torch.allclose(alpha, a, atol=1e-6)
    exp_img = a[..., None] * torch.tensor(color) + (1 - a)[..., None] * torch.tensor(bg)
    assert torch.allclose(img, exp_img, atol=1e-6)


def test_raster_alpha_clamp_and_thresholds():
    # opacity 1 at the centre pixel -> alpha clamped to 0.999
    img, alpha, _, _ = _raster_single(1.0, [1.0, 1.0, 1.0], [0.0, 0.0, 0.0], s=2.0)
    assert abs(float(alpha.max()) - 0.999) < 1e-6
    # opacity below 1/255 never contributes
    img, alpha, _, _ = _raster_single(0.0039, [1.0, 1.0, 1.0], [0.5, 0.5, 0.5])
    assert float(alpha.abs().max()) == 0.0 and torch.allclose(img, torch.full_like(img, 0.5))


def test_raster_depth_order_and_termination():
    fx = fy = 100.0
    size = 16
    vm = _identity_view()
    # two coincident Gaussians, the RED one nearer
    means = torch.tensor([[0.0, 0.0, 2.0], [0.0, 0.0, 1.5]])
    scales = torch.full((2, 3), 0.1)
    quats = torch.tensor([[1.0, 0, 0, 0], [1.0, 0, 0, 0]])
    colors = torch.tensor([[0.0, 0.0, 1.0], [1.0, 0.0, 0.0]])
    opac = torch.tensor([[0.6], [0.5]])
    xys, depths, radii, conics, comp, tiles, _ = R.project_gaussians(means, scales, 1, quats, vm, fx, fy, 8.0, 8.0,
                                                                     size, size, 16)
    img, alpha, aux = R.rasterize_gaussians(xys, depths, radii, conics, tiles, colors, opac, size, size, 16,
                                            background=torch.zeros(3), return_alpha=True, return_aux=True)
    assert aux["gaussian_ids_sorted"].tolist() == [1, 0]          # nearer first
    cy = cx = 8   # pixel centre (8.5, 8.5), Gaussians at (8, 8)
    d2 = 0.5
    a_near = 0.5 * math.exp(-0.5 * d2 * float(conics[1, 0]))
    a_far = 0.6 * math.exp(-0.5 * d2 * float(conics[0, 0]))
    exp = torch.tensor([a_near, 0.0, a_far * (1 - a_near)])
    assert torch.allclose(img[cy, cx], exp, atol=1e-6)
    assert abs(float(alpha[cy, cx]) - (1 - (1 - a_near) * (1 - a_far))) < 1e-6
    # termination: 5 opaque layers; T after k layers = 0.001^k; the 2nd would give 1e-6 <= 1e-4 -> NOT composited
    n = 5
    means = torch.tensor([[0.0, 0.0, 1.0 + 0.1 * i] for i in range(n)])
    scales = torch.full((n, 3), 0.5)
    quats = torch.tensor([[1.0, 0, 0, 0]] * n)
    colors = torch.tensor([[1.0, 0, 0]] + [[0.0, 1.0, 0]] * (n - 1))
    opac = torch.ones(n, 1)
    xys, depths, radii, conics, comp, tiles, _ = R.project_gaussians(means, scales, 1, quats, vm, fx, fy, 8.0, 8.0,
                                                                     size, size, 16)
    img, alpha, aux = R.rasterize_gaussians(xys, depths, radii, conics, tiles, colors, opac, size, size, 16,
                                            background=torch.ones(3), return_alpha=True, return_aux=True)
    assert abs(float(img[8, 8, 0]) - (0.999 + 0.001)) < 1e-6      # red layer + T * white background
    assert abs(float(img[8, 8, 1]) - 0.001) < 1e-6                # the green layers were never composited
    assert abs(float(alpha[8, 8]) - 0.999) < 1e-6
    assert int(aux["final_idx"][8, 8]) == 0


def test_raster_empty_frame_quirk():
    n = 4
    z = torch.zeros
    img, alpha = R.rasterize_gaussians(z(n, 2), z(n), z(n, dtype=torch.int32), z(n, 3), z(n, dtype=torch.int32),
                                       z(n, 3), z(n, 1), 20, 24, 16, background=torch.tensor([0.1, 0.2, 0.3]),
                                       return_alpha=True)
    assert img.shape == (20, 24, 3) and torch.allclose(img[3, 5], torch.tensor([0.1, 0.2, 0.3]))
    assert bool((alpha == 1).all())        # upstream: final_Ts = 0 on the num_intersects < 1 early return
    only = R.rasterize_gaussians(z(n, 2), z(n), z(n, dtype=torch.int32), z(n, 3), z(n, dtype=torch.int32),
                                 z(n, 3), z(n, 1), 20, 24, 16)
    assert isinstance(only, torch.Tensor) and bool((only == 1).all())


def test_viewmat_convention():
    # OpenGL camera at (0,0,3) looking down -z at the origin, y up
    c2w = torch.eye(4)
    c2w[2, 3] = 3.0
    vm = R.viewmat_from_c2w(c2w)
    p = vm @ torch.tensor([0.0, 1.0, 0.0, 1.0])     # a point above the origin
    # OpenCV camera: y down, z forward
    assert torch.allclose(p[:3], torch.tensor([0.0, -1.0, 3.0]))


def test_tile_bins_and_keys_small():
    # three Gaussians with hand-checked tiles on a 32x32 image (2x2 tiles)
    xys = torch.tensor([[8.0, 8.0], [16.0, 16.0], [24.0, 8.0]])
    radii = torch.tensor([2, 3, 2], dtype=torch.int32)
    depths = torch.tensor([3.0, 1.0, 2.0])
    tiles = torch.tensor([1, 4, 1], dtype=torch.int32)
    ni, cum = R.compute_cumulative_intersects(tiles)
    assert ni == 6 and cum.tolist() == [1, 5, 6]
    keys, gids, keys_s, gids_s, bins = R.bin_and_sort_gaussians(3, ni, xys, depths, radii, cum, (2, 2, 1), 16)
    assert gids.tolist() == [0, 1, 1, 1, 1, 2]
    assert (keys >> 32).tolist() == [0, 0, 1, 2, 3, 1]
    one = torch.tensor([1.0]).view(torch.int32).item()
    assert int(keys[1] & 0xFFFFFFFF) == one
    assert gids_s.tolist() == [1, 0, 1, 2, 1, 1]                 # tile 0: depth 1 then 3; tile 1: 1 then 2
    assert bins.tolist() == [[0, 2], [2, 4], [4, 5], [5, 6]]


def test_cov2d_bounds_known_answers():
    """compute_cov2d_bounds: identity covariance -> conic (1,0,1); lambda_max = 1 + sqrt(0.1) (the
    discriminant floor), radius = ceil(3 sqrt(1.316)) = 4; diag(4, 1): conic (0.25, 0, 1), radius
    ceil(3*2) = 6; singular -> zeros."""
    cov = torch.tensor([[1.0, 0.0, 1.0], [4.0, 0.0, 1.0], [1.0, 1.0, 1.0]])
    conic, radius = R.compute_cov2d_bounds(cov)
    assert torch.allclose(conic, torch.tensor([[1.0, 0.0, 1.0], [0.25, 0.0, 1.0], [0.0, 0.0, 0.0]]))
    assert radius.tolist() == [4, 6, 0] and radius.dtype == torch.int32


def test_oracle_rasterizes_any_channel_count():
    """The first three channels of a 5-channel render equal the 3-channel render of the same splats."""
    gs = synth.shell_scene(300, 0, seed=3, scale_mult=0.6)
    cams = synth.orbit_cameras(1, 48, 40)
    k = H.cam_scalars(cams)
    inp = H.view_inputs(gs, cams, 0)
    p = R.project_gaussians(inp["means"], inp["scales"], 1, inp["quats"], inp["viewmat"], k["fx"], k["fy"], k["cx"],
                            k["cy"], k["H"], k["W"], 16)
    cols = torch.rand(300, 5, generator=torch.Generator().manual_seed(1))
    bg = torch.tensor([0.1, 0.2, 0.3, 0.4, 0.5])
    img5, a5 = R.rasterize_gaussians(p[0], p[1], p[2], p[3], p[5], cols, inp["opacities"], k["H"], k["W"], 16,
                                     background=bg, return_alpha=True)
    img3, a3 = R.rasterize_gaussians(p[0], p[1], p[2], p[3], p[5], cols[:, :3].contiguous(), inp["opacities"],
                                     k["H"], k["W"], 16, background=bg[:3], return_alpha=True)
    assert img5.shape == (40, 48, 5) and torch.equal(a5, a3)
    assert torch.allclose(img5[..., :3], img3, atol=1e-6)
