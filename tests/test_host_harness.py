"""The per-Gaussian device math (splatformer_b200/csrc/sb_math.cuh) compiled for the HOST with
g++ -ffp-contract=off (tests/host_harness) against the oracle.  The same source is what the CUDA
projection kernels run with --fmad=false, so these tests pin, without a GPU, (1) the bit-exact
arithmetic contract of the projection / tiling decisions and (2) the hand-derived projection VJP."""
import ctypes

import numpy as np
import pytest
import torch

from host_harness import build as hh_build
from oracle import ref_torch as R
from splatformer_b200 import synth
import helpers as H


@pytest.fixture(scope="module")
def hh():
    return hh_build.load()


def _p(a, ct=ctypes.c_float):
    return a.ctypes.data_as(ctypes.POINTER(ct))


def _project_host(hh, inp, k, bw, clip=0.01):
    n = inp["means"].shape[0]
    m, s, q = (np.ascontiguousarray(inp[x].numpy(), dtype=np.float32) for x in ("means", "scales", "quats"))
    vm = np.ascontiguousarray(inp["viewmat"].numpy(), dtype=np.float32)
    xys = np.zeros((n, 2), np.float32); depths = np.zeros(n, np.float32); radii = np.zeros(n, np.int32)
    conics = np.zeros((n, 3), np.float32); comp = np.zeros(n, np.float32); tiles = np.zeros(n, np.int32)
    cov3d = np.zeros((n, 6), np.float32)
    f = ctypes.c_float
    hh.hh_project_fwd(n, _p(m), _p(s), f(1.0), _p(q), _p(vm), f(k["fx"]), f(k["fy"]), f(k["cx"]), f(k["cy"]),
                      k["H"], k["W"], bw, f(clip), _p(xys), _p(depths), _p(radii, ctypes.c_int), _p(conics),
                      _p(comp), _p(tiles, ctypes.c_int), _p(cov3d))
    return xys, depths, radii, conics, comp, tiles, cov3d


@pytest.mark.parametrize("n,size,bw,mult,seed", [(5000, (128, 128), 16, 0.15, 1), (3000, (200, 120), 16, 0.3, 2),
                                                 (2000, (96, 80), 12, 0.3, 3), (20000, (512, 512), 16, 0.15, 4)])
def test_projection_forward_bit_exact(hh, n, size, bw, mult, seed):
    gs = synth.shell_scene(n, 0, seed=seed, scale_mult=mult)
    cams = synth.orbit_cameras(3, size[0], size[1])
    k = H.cam_scalars(cams)
    for view in range(3):
        inp = H.view_inputs(gs, cams, view)
        ref = R.project_gaussians(inp["means"], inp["scales"], 1, inp["quats"], inp["viewmat"], k["fx"], k["fy"],
                                  k["cx"], k["cy"], k["H"], k["W"], bw)
        got = _project_host(hh, inp, k, bw)
        assert int((ref[2] > 0).sum()) > n // 10
        for name, a, b in zip(("xys", "depths", "radii", "conics", "compensation", "num_tiles_hit", "cov3d"), got, ref):
            assert np.array_equal(a, b.numpy()), f"{name}: host build of sb_math.cuh differs from the oracle"


def test_projection_degenerate_inputs_cull_identically(hh):
    gs = synth.shell_scene(64, 0, seed=5, scale_mult=0.6)
    gs["means"][0] = torch.tensor([float("nan"), 0.5, 0.5])
    gs["means"][1] = torch.tensor([1e30, 0.5, 0.5])
    gs["scales"][2] = torch.tensor([80.0, 80.0, 80.0])            # exp overflows to inf
    gs["scales"][3] = torch.tensor([-80.0, -80.0, -80.0])         # exp underflows to 0
    gs["quats"][4] = torch.tensor([0.0, 0.0, 0.0, 0.0])
    cams = synth.orbit_cameras(1, 64, 64)
    k = H.cam_scalars(cams)
    inp = H.view_inputs(gs, cams, 0)
    ref = R.project_gaussians(inp["means"], inp["scales"], 1, inp["quats"], inp["viewmat"], k["fx"], k["fy"],
                              k["cx"], k["cy"], k["H"], k["W"], 16)
    got = _project_host(hh, inp, k, 16)
    assert np.array_equal(got[2], ref[2].numpy()) and np.array_equal(got[5], ref[5].numpy())
    for i in (0, 1, 2, 4):
        assert got[2][i] == 0 and got[5][i] == 0


def test_projection_backward_matches_autograd(hh):
    n = 800
    gs = synth.shell_scene(n, 0, seed=6, scale_mult=0.6)
    cams = synth.orbit_cameras(2, 96, 72)
    k = H.cam_scalars(cams)
    inp = H.view_inputs(gs, cams, 1, dtype=torch.float64)
    leaves = [inp[x].clone().requires_grad_() for x in ("means", "scales", "quats", "viewmat")]
    ref = R.project_gaussians(leaves[0], leaves[1], 1, leaves[2], leaves[3], k["fx"], k["fy"], k["cx"], k["cy"],
                              k["H"], k["W"], 16)
    torch.manual_seed(0)
    ws = {i: torch.randn(ref[i].shape, dtype=torch.float64) for i in (0, 1, 3, 4, 6)}
    sum((ref[i] * w).sum() for i, w in ws.items()).backward()
    f32 = lambda t: np.ascontiguousarray(t.detach().numpy(), dtype=np.float32)
    m, s, q, vm = (f32(inp[x]) for x in ("means", "scales", "quats", "viewmat"))
    vx, vd, vc, vcomp, vc3 = (f32(ws[i]) for i in (0, 1, 3, 4, 6))
    v_m = np.zeros((n, 3), np.float32); v_s = np.zeros((n, 3), np.float32); v_q = np.zeros((n, 4), np.float32)
    v_vm = np.zeros(12, np.float32)
    f = ctypes.c_float
    hh.hh_project_bwd(n, _p(m), _p(s), f(1.0), _p(q), _p(vm), f(k["fx"]), f(k["fy"]), f(k["cx"]), f(k["cy"]),
                      k["H"], k["W"], 16, f(0.01), _p(vx), _p(vd), _p(vc), _p(vcomp), _p(vc3), _p(v_m), _p(v_s),
                      _p(v_q), _p(v_vm))
    H.assert_grad_close(torch.from_numpy(v_m), leaves[0].grad, "v_means")
    H.assert_grad_close(torch.from_numpy(v_s), leaves[1].grad, "v_scales")
    H.assert_grad_close(torch.from_numpy(v_q), leaves[2].grad, "v_quats")
    H.assert_grad_close(torch.from_numpy(v_vm).reshape(3, 4), leaves[3].grad[:3], "v_viewmat")


def test_sh_basis_matches_oracle(hh):
    torch.manual_seed(1)
    d = torch.randn(500, 3)
    for deg in range(5):
        K = (deg + 1) ** 2
        Y = np.zeros((500, K), np.float32)
        dn = np.ascontiguousarray(d.numpy())
        hh.hh_sh_basis(500, deg, _p(dn), _p(Y))
        assert np.allclose(Y, R.sh_basis(deg, d.double()).numpy(), atol=2e-6)


def test_visible_extent_is_conservative(hh):
    """The raster kernels skip a Gaussian for a warp when its alpha >= 1/255 region misses the warp's
    pixels; the extent must therefore contain every point where alpha >= 1/255."""
    rng = np.random.default_rng(0)
    n = 4000
    a = rng.uniform(0.002, 3.0, n).astype(np.float32)
    c = rng.uniform(0.002, 3.0, n).astype(np.float32)
    b = (rng.uniform(-0.95, 0.95, n) * np.sqrt(a * c)).astype(np.float32)
    op = rng.uniform(0.0, 1.0, n).astype(np.float32)
    conics = np.ascontiguousarray(np.stack([a, b, c], 1))
    hx = np.zeros(n, np.float32); hy = np.zeros(n, np.float32)
    hh.hh_visible_extent(n, _p(conics), _p(op), _p(hx), _p(hy))
    never = op * 255.0 < 1.0
    assert np.all(hx[never] < 0) and np.all(hy[never] < 0)
    # sample points on the alpha = 1/255 contour: sigma = ln(255 op)
    vis = ~never
    tau = np.log(255.0 * op[vis].astype(np.float64))
    A, B, C = (x[vis].astype(np.float64) for x in (a, b, c))
    for theta in np.linspace(0, 2 * np.pi, 73):
        ux, uy = np.cos(theta), np.sin(theta)
        q = 0.5 * (A * ux * ux + C * uy * uy) + B * ux * uy       # sigma along the ray = q r^2
        r = np.sqrt(tau / q)
        assert np.all(np.abs(r * ux) <= hx[vis] + 1e-6) and np.all(np.abs(r * uy) <= hy[vis] + 1e-6)
    # not absurdly loose either: within 1 % + 0.01 px of the true extreme
    det = A * C - B * B
    assert np.all(hx[vis] <= np.sqrt(2 * tau * C / det) * 1.01 + 0.01)
