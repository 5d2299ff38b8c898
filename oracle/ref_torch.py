"""ORACLE (test infrastructure, not product code) -- parity unpinned.

Vectorised pure-PyTorch restatement of the differentiable splat render that
SplatFormer reaches through the gsplat v0.1.11 Python API
(`/root/reference/utils/gs_utils.py:78` spherical_harmonics, `:82-95`
project_gaussians, `:96-109` rasterize_gaussians).  gsplat itself is an
un-vendored pip-from-git dependency (`/root/reference/README.md:27`) whose
source is not on this machine, so this file restates its published algorithm
 and anchors on the reference's call sites.  There are
no reference-owned golden vectors for this path: **parity unpinned**; the pins
are the analytic known-answer tests in `tests/test_oracle_known_answers.py`,
the literal per-pixel loop in `oracle/ref_loop.py`, float64 finite differences
and the frozen fixtures under `tests/golden/`.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU-baseline /
`--impl reference` legs may import this module.  The product path
(`splatformer_b200`) never does and has no CPU fallback.

Arithmetic contract (what "bit-exact tiling" means).  Every float operation in
`project_gaussians` below is a separately rounded IEEE-754 binary32 add, mul,
div or sqrt, evaluated in exactly the order written (Python's left-to-right
association, no fused multiply-add, no matmul).  The CUDA projection kernel is
compiled with `--fmad=false` and repeats the same sequence, so `xys`, `depths`,
`conics`, `radii`, `num_tiles_hit`, the 64-bit intersection keys, the sorted
Gaussian ids and `tile_bins` agree bit for bit between this file on a CPU and
the kernels on a B200.  Integer decisions are always taken in float32, also
when the differentiable math runs in float64 (`dtype=torch.float64` inputs).
"""
from __future__ import annotations

import math
from typing import Optional, Tuple

import torch
from torch import Tensor

# ----------------------------------------------------------------------------
# A.1 spherical harmonics
# ----------------------------------------------------------------------------
SH_C0 = 0.28209479177387814
SH_C1 = 0.4886025119029199
SH_C2 = (1.0925484305920792, -1.0925484305920792, 0.31539156525252005,
         -1.0925484305920792, 0.5462742152960396)
SH_C3 = (-0.5900435899266435, 2.890611442640554, -0.4570457994644658,
         0.3731763325901154, -0.4570457994644658, 1.445305721320277,
         -0.5900435899266435)
SH_C4 = (2.5033429417967046, -1.7701307697799304, 0.9461746957575601,
         -0.6690465435572892, 0.10578554691520431, -0.6690465435572892,
         0.47308734787878004, -1.7701307697799304, 0.6258357354491761)


def num_sh_bases(degree: int) -> int:
    if degree < 0 or degree > 4:
        raise ValueError(f"SH degree must be in 0..4, got {degree}")
    return (degree + 1) ** 2


def deg_from_sh(num_bases: int) -> int:
    for d in range(5):
        if (d + 1) ** 2 == num_bases:
            return d
    raise ValueError(f"invalid number of SH bases: {num_bases}")


def sh_basis(degree: int, dirs: Tensor) -> Tensor:
    """Real SH basis values Y_k(dir) for k < (degree+1)^2, shape (N, K_used).

    The direction is normalised here as upstream does inside its kernel
    (SURVEY.md A.1); SplatFormer already passes unit vectors
    (`/root/reference/utils/gs_utils.py:67-69`).
    """
    n = torch.sqrt(dirs[:, 0] * dirs[:, 0] + dirs[:, 1] * dirs[:, 1] + dirs[:, 2] * dirs[:, 2])
    x, y, z = dirs[:, 0] / n, dirs[:, 1] / n, dirs[:, 2] / n
    out = [torch.full_like(x, SH_C0)]
    if degree >= 1:
        out += [-SH_C1 * y, SH_C1 * z, -SH_C1 * x]
    if degree >= 2:
        xx, yy, zz = x * x, y * y, z * z
        xy, yz, xz = x * y, y * z, x * z
        out += [SH_C2[0] * xy, SH_C2[1] * yz, SH_C2[2] * (2.0 * zz - xx - yy),
                SH_C2[3] * xz, SH_C2[4] * (xx - yy)]
    if degree >= 3:
        out += [SH_C3[0] * y * (3.0 * xx - yy), SH_C3[1] * xy * z,
                SH_C3[2] * y * (4.0 * zz - xx - yy),
                SH_C3[3] * z * (2.0 * zz - 3.0 * xx - 3.0 * yy),
                SH_C3[4] * x * (4.0 * zz - xx - yy), SH_C3[5] * z * (xx - yy),
                SH_C3[6] * x * (xx - 3.0 * yy)]
    if degree >= 4:
        out += [SH_C4[0] * xy * (xx - yy), SH_C4[1] * yz * (3.0 * xx - yy),
                SH_C4[2] * xy * (7.0 * zz - 1.0), SH_C4[3] * yz * (7.0 * zz - 3.0),
                SH_C4[4] * (zz * (35.0 * zz - 30.0) + 3.0),
                SH_C4[5] * xz * (7.0 * zz - 3.0), SH_C4[6] * (xx - yy) * (7.0 * zz - 1.0),
                SH_C4[7] * xz * (xx - 3.0 * yy),
                SH_C4[8] * (xx * (xx - 3.0 * yy) - yy * (3.0 * xx - yy))]
    return torch.stack(out, dim=1)


def spherical_harmonics(degrees_to_use: int, viewdirs: Tensor, coeffs: Tensor,
                        method: str = "poly") -> Tensor:
    """gsplat.spherical_harmonics (call site gs_utils.py:78).

    rgb[n,c] = sum_{k<(degrees_to_use+1)^2} Y_k(viewdirs[n]) * coeffs[n,k,c];
    no +0.5 and no clamp (the wrapper adds them, gs_utils.py:79).  Gradient
    flows to `coeffs` only (viewdirs is treated as a constant, as upstream).
    """
    if method not in ("poly", "fast"):
        raise ValueError(f"unknown SH method {method!r}")
    if coeffs.dim() != 3 or coeffs.shape[2] != 3:
        raise ValueError("coeffs must be (N, K, 3)")
    storage_deg = deg_from_sh(coeffs.shape[1])
    if degrees_to_use > storage_deg:
        raise AssertionError("coeffs must have at least (degrees_to_use+1)^2 bases")
    basis = sh_basis(degrees_to_use, viewdirs.detach().to(coeffs.dtype))
    k_used = basis.shape[1]
    return (basis[:, :, None] * coeffs[:, :k_used, :]).sum(dim=1)


# ----------------------------------------------------------------------------
# A.2 / A.3 projection
# ----------------------------------------------------------------------------
RADIUS_MAX = 1073741824.0  # 2^30


def _f32(v) -> float:
    """Round a Python scalar to binary32 (what a `float` kernel argument holds)."""
    return float(torch.tensor(float(v), dtype=torch.float32))


def _sqrt(x: Tensor) -> Tensor:
    """Correctly rounded square root.  torch's CPU float32 sqrt goes through MKL VML and is
    off by one ulp for ~0.7 % of inputs; sqrt in binary64 followed by one rounding to binary32
    is the IEEE result (53 >= 2*24+2), which is what CUDA's sqrtf returns."""
    if x.dtype == torch.float32:
        return torch.sqrt(x.double()).float()
    return torch.sqrt(x)


def _tile_bbox(cx_pix: Tensor, cy_pix: Tensor, radius: Tensor, tiles_x: int, tiles_y: int,
               block_width: int):
    """A.3: inclusive min / exclusive max tile indices, float32 then trunc."""
    bw = float(block_width)
    tcx, tcy, tr = cx_pix / bw, cy_pix / bw, radius / bw
    def clampi(v: Tensor, hi: int) -> Tensor:
        # (int) cast truncates toward zero; values are far inside int32 range
        # for any pixel coordinate that survives the tile_area test, but guard
        # against overflow of astronomically large coordinates first.
        v = torch.clamp(v, min=-2.0e9, max=2.0e9)
        return torch.clamp(v.to(torch.int64), min=0, max=hi)
    min_x = clampi(tcx - tr, tiles_x)
    max_x = clampi(tcx + tr + 1.0, tiles_x)
    min_y = clampi(tcy - tr, tiles_y)
    max_y = clampi(tcy + tr + 1.0, tiles_y)
    return min_x, max_x, min_y, max_y


def project_gaussians(means3d: Tensor, scales: Tensor, glob_scale: float, quats: Tensor,
                      viewmat: Tensor, fx: float, fy: float, cx: float, cy: float,
                      img_height: int, img_width: int, block_width: int,
                      clip_thresh: float = 0.01):
    """gsplat.project_gaussians (call site gs_utils.py:82-95), 7-tuple return.

    Returns (xys, depths, radii, conics, compensation, num_tiles_hit, cov3d).
    Culled Gaussians have radii = num_tiles_hit = 0 and zeros elsewhere
    (cov3d is written once the near-plane test passed, A.2).
    """
    assert 2 <= block_width <= 16, "block_width must be between 2 and 16"
    dt = means3d.dtype
    # intrinsics reach the kernels as binary32 scalars; round them first so the
    # float64 mode differentiates exactly the same function
    fx, fy, cx, cy = _f32(fx), _f32(fy), _f32(cx), _f32(cy)
    clip_thresh = _f32(clip_thresh)
    vm = viewmat.to(dt)
    w = [[vm[i, j] for j in range(3)] for i in range(3)]
    p = [vm[i, 3] for i in range(3)]
    x, y, z = means3d[:, 0], means3d[:, 1], means3d[:, 2]

    tx = w[0][0] * x + w[0][1] * y + w[0][2] * z + p[0]
    ty = w[1][0] * x + w[1][1] * y + w[1][2] * z + p[1]
    tz = w[2][0] * x + w[2][1] * y + w[2][2] * z + p[2]
    front = tz.detach() > clip_thresh
    # safe denominators for the culled rows (double-where keeps autograd finite)
    tz_s = torch.where(front, tz, torch.ones_like(tz))

    # --- cov3d = (R S)(R S)^T ------------------------------------------------
    qn = _sqrt(quats[:, 0] * quats[:, 0] + quats[:, 1] * quats[:, 1]
                    + quats[:, 2] * quats[:, 2] + quats[:, 3] * quats[:, 3])
    qw, qx, qy, qz = quats[:, 0] / qn, quats[:, 1] / qn, quats[:, 2] / qn, quats[:, 3] / qn
    r = [[1.0 - 2.0 * (qy * qy + qz * qz), 2.0 * (qx * qy - qw * qz), 2.0 * (qx * qz + qw * qy)],
         [2.0 * (qx * qy + qw * qz), 1.0 - 2.0 * (qx * qx + qz * qz), 2.0 * (qy * qz - qw * qx)],
         [2.0 * (qx * qz - qw * qy), 2.0 * (qy * qz + qw * qx), 1.0 - 2.0 * (qx * qx + qy * qy)]]
    s = [glob_scale * scales[:, j] for j in range(3)]
    m = [[r[i][j] * s[j] for j in range(3)] for i in range(3)]
    def sig(i, j):
        return m[i][0] * m[j][0] + m[i][1] * m[j][1] + m[i][2] * m[j][2]
    s00, s01, s02, s11, s12, s22 = sig(0, 0), sig(0, 1), sig(0, 2), sig(1, 1), sig(1, 2), sig(2, 2)
    sg = [[s00, s01, s02], [s01, s11, s12], [s02, s12, s22]]

    # --- EWA projection -----------------------------------------------------
    # limits are computed in float32 regardless of dt so that both precisions clamp alike:
    # lim = 1.3f * ((0.5f * size) / f)
    f32 = torch.float32
    lim_x = float(torch.tensor(1.3, dtype=f32) * (torch.tensor(0.5 * img_width, dtype=f32) / torch.tensor(fx, dtype=f32)))
    lim_y = float(torch.tensor(1.3, dtype=f32) * (torch.tensor(0.5 * img_height, dtype=f32) / torch.tensor(fy, dtype=f32)))
    rz = 1.0 / tz_s
    rz2 = rz * rz
    txc = tz_s * torch.clamp(tx * rz, min=-lim_x, max=lim_x)
    tyc = tz_s * torch.clamp(ty * rz, min=-lim_y, max=lim_y)
    j00 = fx * rz
    j02 = -(fx * txc) * rz2
    j11 = fy * rz
    j12 = -(fy * tyc) * rz2
    t0 = [j00 * w[0][j] + j02 * w[2][j] for j in range(3)]
    t1 = [j11 * w[1][j] + j12 * w[2][j] for j in range(3)]
    a0 = [t0[0] * sg[0][j] + t0[1] * sg[1][j] + t0[2] * sg[2][j] for j in range(3)]
    a1 = [t1[0] * sg[0][j] + t1[1] * sg[1][j] + t1[2] * sg[2][j] for j in range(3)]
    c00 = a0[0] * t0[0] + a0[1] * t0[1] + a0[2] * t0[2]
    c01 = a0[0] * t1[0] + a0[1] * t1[1] + a0[2] * t1[2]
    c11 = a1[0] * t1[0] + a1[1] * t1[1] + a1[2] * t1[2]

    det_orig = c00 * c11 - c01 * c01
    ca = c00 + 0.3
    cb = c01
    cc = c11 + 0.3
    det = ca * cc - cb * cb
    det_ok = det.detach() != 0
    det_s = torch.where(det_ok, det, torch.ones_like(det))
    ratio = det_orig / det_s
    pos = ratio.detach() > 0
    comp = torch.where(pos, _sqrt(torch.where(pos, ratio, torch.ones_like(ratio))),
                       torch.zeros_like(ratio))
    inv_det = 1.0 / det_s
    conic = torch.stack([cc * inv_det, -cb * inv_det, ca * inv_det], dim=1)

    bmid = 0.5 * (ca + cc)
    disc = _sqrt(torch.clamp(bmid * bmid - det_s, min=0.1))
    v1 = bmid + disc
    v2 = bmid - disc
    radius_f = torch.ceil(3.0 * _sqrt(torch.maximum(v1, v2))).detach()

    rw = 1.0 / (tz_s + 1e-6)
    px = fx * (tx * rw) + cx
    py = fy * (ty * rw) + cy

    tiles_x = (img_width + block_width - 1) // block_width
    tiles_y = (img_height + block_width - 1) // block_width
    # integer decisions always in float32
    if dt == torch.float32:
        pxf, pyf, rf = px.detach(), py.detach(), radius_f
    else:
        # float64 mode is only used with rasterize-level tests; callers that need
        # bit-exact tiling use float32 inputs.
        pxf, pyf, rf = px.detach().float(), py.detach().float(), radius_f.float()
    min_x, max_x, min_y, max_y = _tile_bbox(pxf, pyf, rf, tiles_x, tiles_y, block_width)
    tile_area = (max_x - min_x) * (max_y - min_y)
    # a NaN/inf anywhere in the chain (degenerate input) culls the Gaussian
    # (radius capped at 2^30 so that the int32 store and the bbox arithmetic stay exact)
    finite = torch.isfinite(radius_f) & (radius_f <= RADIUS_MAX) & torch.isfinite(px.detach()) \
        & torch.isfinite(py.detach())
    valid = front & det_ok & (tile_area > 0) & finite

    zero = torch.zeros_like(tz)
    xys = torch.stack([torch.where(valid, px, zero), torch.where(valid, py, zero)], dim=1)
    depths = torch.where(valid, tz, zero)
    radii = torch.where(valid, radius_f, torch.zeros_like(radius_f)).to(torch.int32)
    conics = torch.where(valid[:, None], conic, torch.zeros_like(conic))
    compensation = torch.where(valid, comp, zero)
    num_tiles_hit = torch.where(valid, tile_area, torch.zeros_like(tile_area)).to(torch.int32)
    cov3d = torch.stack([s00, s01, s02, s11, s12, s22], dim=1)
    cov3d = torch.where(front[:, None], cov3d, torch.zeros_like(cov3d))
    return xys, depths, radii, conics, compensation, num_tiles_hit, cov3d


# ----------------------------------------------------------------------------
# A.4 intersections, sort, bins
# ----------------------------------------------------------------------------
def compute_cov2d_bounds(cov2d: Tensor) -> Tuple[Tensor, Tensor]:
    """gsplat.utils.compute_cov2d_bounds [upstream]: conic = inverse of (a b; b c), radius =
    ceil(3 sqrt(lambda_max)) with the discriminant floored at 0.1 (same lines as A.2)."""
    a, b, c = cov2d[:, 0], cov2d[:, 1], cov2d[:, 2]
    det = a * c - b * b
    ok = det != 0
    det_s = torch.where(ok, det, torch.ones_like(det))
    inv = 1.0 / det_s
    conic = torch.stack([c * inv, -b * inv, a * inv], dim=1)
    mid = 0.5 * (a + c)
    disc = _sqrt(torch.clamp(mid * mid - det_s, min=0.1))
    radius = torch.ceil(3.0 * _sqrt(torch.maximum(mid + disc, mid - disc)))
    conic = torch.where(ok[:, None], conic, torch.zeros_like(conic))
    radius = torch.where(ok, radius, torch.zeros_like(radius))
    return conic, radius.to(torch.int32)


def compute_cumulative_intersects(num_tiles_hit: Tensor) -> Tuple[int, Tensor]:
    cum = torch.cumsum(num_tiles_hit, dim=0, dtype=torch.int32)
    return (int(cum[-1].item()) if cum.numel() else 0), cum


def map_gaussian_to_intersects(num_points: int, num_intersects: int, xys: Tensor, depths: Tensor,
                               radii: Tensor, cum_tiles_hit: Tensor, tile_bounds, block_width: int):
    """A.4 key emission: key = tile_id << 32 | float32 bits of depth."""
    tiles_x, tiles_y = int(tile_bounds[0]), int(tile_bounds[1])
    xf = xys.detach().float()
    min_x, max_x, min_y, max_y = _tile_bbox(xf[:, 0], xf[:, 1], radii.float(), tiles_x, tiles_y,
                                            block_width)
    hit = radii > 0
    wx = torch.where(hit, max_x - min_x, torch.zeros_like(min_x))
    wy = torch.where(hit, max_y - min_y, torch.zeros_like(min_y))
    cnt = wx * wy
    start = cum_tiles_hit.to(torch.int64) - cnt  # == (g==0 ? 0 : cum[g-1]) when cnt == num_tiles_hit
    gid = torch.repeat_interleave(torch.arange(num_points, dtype=torch.int64), cnt)
    assert gid.numel() == num_intersects, (gid.numel(), num_intersects)
    local = torch.arange(num_intersects, dtype=torch.int64) - start[gid]
    ty = min_y[gid] + local // torch.clamp(wx[gid], min=1)
    tx = min_x[gid] + local % torch.clamp(wx[gid], min=1)
    tile_id = ty * tiles_x + tx
    depth_bits = depths.detach().float().contiguous().view(torch.int32).to(torch.int64) & 0xFFFFFFFF
    isect_ids = (tile_id << 32) | depth_bits[gid]
    return isect_ids, gid.to(torch.int32)


def get_tile_bin_edges(num_intersects: int, isect_ids_sorted: Tensor, tile_bounds) -> Tensor:
    tiles = int(tile_bounds[0]) * int(tile_bounds[1])
    bins = torch.zeros(tiles, 2, dtype=torch.int32)
    if num_intersects == 0:
        return bins
    tile = (isect_ids_sorted >> 32).to(torch.int64)
    idx = torch.arange(num_intersects, dtype=torch.int32)
    first = torch.ones(num_intersects, dtype=torch.bool)
    first[1:] = tile[1:] != tile[:-1]
    last = torch.ones(num_intersects, dtype=torch.bool)
    last[:-1] = first[1:]
    bins[tile[first], 0] = idx[first]
    bins[tile[last], 1] = idx[last] + 1
    return bins


def bin_and_sort_gaussians(num_points: int, num_intersects: int, xys: Tensor, depths: Tensor,
                           radii: Tensor, cum_tiles_hit: Tensor, tile_bounds, block_width: int):
    isect_ids, gaussian_ids = map_gaussian_to_intersects(
        num_points, num_intersects, xys, depths, radii, cum_tiles_hit, tile_bounds, block_width)
    isect_ids_sorted, order = torch.sort(isect_ids, stable=True)
    gaussian_ids_sorted = torch.gather(gaussian_ids, 0, order)
    tile_bins = get_tile_bin_edges(num_intersects, isect_ids_sorted, tile_bounds)
    return isect_ids, gaussian_ids, isect_ids_sorted, gaussian_ids_sorted, tile_bins


def tile_lists(xys: Tensor, depths: Tensor, radii: Tensor, tile_bounds, block_width: int, tiles):
    """The sorted Gaussian-id list of each tile in `tiles`, computed tile by tile instead of through the
    global sort: the Gaussians whose tile bbox (A.3) contains the tile, ordered by (depth bits, Gaussian
    id) -- which is what the stable sort of A.4 leaves in that tile's bin (keys of one tile differ only in
    the depth bits; ties keep emission order = ascending id).  For frames whose full key array is too
    large for this CPU code (BASELINE configs 4 and 5); tests/test_oracle_crosscheck.py checks it against
    bin_and_sort_gaussians.  Returns (gaussian_ids_sorted, tile_bins) in the layout rasterize_forward
    takes, holding only the requested tiles."""
    tiles_x, tiles_y = int(tile_bounds[0]), int(tile_bounds[1])
    xf = xys.detach().float()
    min_x, max_x, min_y, max_y = _tile_bbox(xf[:, 0], xf[:, 1], radii.float(), tiles_x, tiles_y, block_width)
    hit = radii > 0
    depth_bits = depths.detach().float().contiguous().view(torch.int32).to(torch.int64) & 0xFFFFFFFF
    bins = torch.zeros(tiles_x * tiles_y, 2, dtype=torch.int32)
    parts, off = [], 0
    for t in sorted(set(int(t) for t in tiles)):
        ty, tx = divmod(t, tiles_x)
        ids = torch.nonzero(hit & (min_x <= tx) & (tx < max_x) & (min_y <= ty) & (ty < max_y)).flatten()
        if ids.numel() == 0:
            continue
        order = torch.sort((depth_bits[ids] << 32) | ids, stable=True).indices
        parts.append(ids[order].to(torch.int32))
        bins[t, 0], bins[t, 1] = off, off + ids.numel()
        off += ids.numel()
    gids = torch.cat(parts) if parts else torch.zeros(0, dtype=torch.int32)
    return gids, bins


# ----------------------------------------------------------------------------
# A.5 rasterise forward (autograd supplies A.6)
# ----------------------------------------------------------------------------
ALPHA_MIN = 1.0 / 255.0
ALPHA_MAX = 0.999
T_MIN = 1e-4


def _composite_tile(dx: Tensor, dy: Tensor, conic: Tensor, opac: Tensor, rgb: Tensor):
    """dx, dy: (P, n) = xys[g] - pixel centre.  Returns pix (P,3), T_final (P,),
    last kept list position (P,) (-1 if none) and the decision margin (P,)."""
    sigma = 0.5 * (conic[None, :, 0] * dx * dx + conic[None, :, 2] * dy * dy) + conic[None, :, 1] * dx * dy
    alpha_raw = opac[None, :] * torch.exp(-sigma)
    alpha = torch.clamp(alpha_raw, max=ALPHA_MAX)
    valid = (sigma.detach() >= 0) & (alpha.detach() >= ALPHA_MIN)
    a = torch.where(valid, alpha, torch.zeros_like(alpha))
    one_minus = 1.0 - a
    t_incl = torch.cumprod(one_minus, dim=1)
    t_excl = torch.cat([torch.ones_like(t_incl[:, :1]), t_incl[:, :-1]], dim=1)
    stop = valid & (t_incl.detach() <= T_MIN)
    stopped = torch.cumsum(stop.to(torch.int32), dim=1) > 0        # includes the stopping Gaussian
    keep = valid & ~stopped
    wgt = torch.where(keep, a * t_excl, torch.zeros_like(a))
    pix = wgt @ rgb
    t_final = torch.prod(torch.where(keep, one_minus, torch.ones_like(one_minus)), dim=1)
    n = dx.shape[1]
    pos = torch.arange(n, dtype=torch.int64)
    last = torch.where(keep, pos[None, :], torch.full_like(pos, -1)[None, :]).amax(dim=1) if n else \
        torch.full((dx.shape[0],), -1, dtype=torch.int64)
    # decision margins: relative distance of every *reached* decision from its threshold
    reached = ~(torch.cumsum(stop.to(torch.int32), dim=1) - stop.to(torch.int32) > 0)
    m_alpha = torch.abs(alpha_raw.detach() / ALPHA_MIN - 1.0)
    m_t = torch.where(valid, torch.abs(t_incl.detach() / T_MIN - 1.0), torch.full_like(m_alpha, 1e9))
    m_clamp = torch.abs(alpha_raw.detach() / ALPHA_MAX - 1.0)
    m = torch.minimum(torch.minimum(m_alpha, m_t), m_clamp)
    m = torch.where(reached, m, torch.full_like(m, 1e9))
    margin = m.amin(dim=1) if n else torch.full((dx.shape[0],), 1e9, dtype=dx.dtype)
    return pix, t_final, last, margin


def rasterize_forward(xys: Tensor, conics: Tensor, colors: Tensor, opacity: Tensor,
                      gaussian_ids_sorted: Tensor, tile_bins: Tensor, img_height: int,
                      img_width: int, block_width: int, background: Tensor, tiles=None):
    """Per-tile front-to-back compositing (A.5).  Differentiable in xys, conics,
    colors, opacity.  Returns out_img (H,W,3), final_Ts (H,W), final_idx (H,W) int32
    and the per-pixel decision margin (H,W) used by tests to set aside pixels whose
    alpha>=1/255 / T<=1e-4 / 0.999-clamp decision lies within rounding noise.
    `tiles` (iterable of tile indices) restricts the compositing to those tiles -- tiles are
    independent, so a subset of a frame too large for this CPU code is still an exact check of
    the pixels it covers; all other pixels are left at T = 1 (background)."""
    dt = colors.dtype
    tiles_x = (img_width + block_width - 1) // block_width
    tiles_y = (img_height + block_width - 1) // block_width
    opac = opacity.reshape(-1)
    channels = colors.shape[-1]                     # 3 for SplatFormer; N-channel variant = upstream nd_rasterize
    img = torch.zeros(img_height, img_width, channels, dtype=dt)
    t_img = torch.ones(img_height, img_width, dtype=dt)
    idx_img = torch.zeros(img_height, img_width, dtype=torch.int32)
    margin_img = torch.full((img_height, img_width), 1e9, dtype=dt)
    img_parts, t_parts = [], []
    bins = tile_bins.to(torch.int64)
    occupied = torch.nonzero(bins[:, 1] > bins[:, 0]).flatten().tolist()
    if tiles is not None:
        wanted = set(int(t) for t in tiles)
        occupied = [t for t in occupied if t in wanted]
    for t in occupied:
        start, end = int(bins[t, 0]), int(bins[t, 1])
        ids = gaussian_ids_sorted[start:end].to(torch.int64)
        ty, tx = divmod(t, tiles_x)
        y0, x0 = ty * block_width, tx * block_width
        y1, x1 = min(y0 + block_width, img_height), min(x0 + block_width, img_width)
        pyv = torch.arange(y0, y1, dtype=dt) + 0.5
        pxv = torch.arange(x0, x1, dtype=dt) + 0.5
        gy, gx = torch.meshgrid(pyv, pxv, indexing="ij")
        gx, gy = gx.reshape(-1), gy.reshape(-1)
        g_xy = xys[ids]
        dx = g_xy[None, :, 0] - gx[:, None]
        dy = g_xy[None, :, 1] - gy[:, None]
        pix, t_fin, last, margin = _composite_tile(dx, dy, conics[ids], opac[ids], colors[ids])
        img_parts.append((y0, y1, x0, x1, pix))
        t_parts.append(t_fin)
        idx_img[y0:y1, x0:x1] = torch.where(last >= 0, last + start, torch.zeros_like(last)).to(
            torch.int32).reshape(y1 - y0, x1 - x0)
        margin_img[y0:y1, x0:x1] = margin.reshape(y1 - y0, x1 - x0)
    # assemble differentiably (no in-place writes into leaf-dependent tensors)
    if img_parts:
        flat_index, flat_pix, flat_t = [], [], []
        for (y0, y1, x0, x1, pix), t_fin in zip(img_parts, t_parts):
            yy, xx = torch.meshgrid(torch.arange(y0, y1), torch.arange(x0, x1), indexing="ij")
            flat_index.append((yy * img_width + xx).reshape(-1))
            flat_pix.append(pix)
            flat_t.append(t_fin)
        index = torch.cat(flat_index)
        img = img.reshape(-1, channels).index_put((index,), torch.cat(flat_pix)).reshape(
            img_height, img_width, channels)
        t_img = t_img.reshape(-1).index_put((index,), torch.cat(flat_t)).reshape(img_height, img_width)
    out_img = img + t_img[:, :, None] * background.to(dt)[None, None, :]
    return out_img, t_img, idx_img, margin_img


def rasterize_gaussians(xys: Tensor, depths: Tensor, radii: Tensor, conics: Tensor,
                        num_tiles_hit: Tensor, colors: Tensor, opacity: Tensor, img_height: int,
                        img_width: int, block_width: int, background: Optional[Tensor] = None,
                        return_alpha: bool = False, return_aux: bool = False, tiles=None):
    """gsplat.rasterize_gaussians (call site gs_utils.py:96-109).  `tiles`: see rasterize_forward."""
    assert 2 <= block_width <= 16, "block_width must be between 2 and 16"
    if colors.dtype == torch.uint8:
        colors = colors.float() / 255
    if background is not None:
        assert background.shape[0] == colors.shape[-1], "incorrect shape of background color tensor"
    else:
        background = torch.ones(colors.shape[-1], dtype=colors.dtype)
    if xys.dim() != 2 or xys.shape[1] != 2:
        raise ValueError("xys must have dimensions (N, 2)")
    if colors.dim() != 2:
        raise ValueError("colors must have dimensions (N, D)")
    tiles_x = (img_width + block_width - 1) // block_width
    tiles_y = (img_height + block_width - 1) // block_width
    num_points = xys.shape[0]
    num_intersects, cum = compute_cumulative_intersects(num_tiles_hit)
    dt = colors.dtype
    if num_intersects < 1:
        out_img = torch.ones(img_height, img_width, colors.shape[-1], dtype=dt) * background.to(dt)
        final_ts = torch.zeros(img_height, img_width, dtype=dt)
        aux = dict(num_intersects=0, final_idx=torch.zeros(img_height, img_width, dtype=torch.int32),
                   margin=torch.full((img_height, img_width), 1e9, dtype=dt),
                   gaussian_ids_sorted=torch.zeros(0, dtype=torch.int32),
                   isect_ids_sorted=torch.zeros(0, dtype=torch.int64),
                   tile_bins=torch.zeros(tiles_x * tiles_y, 2, dtype=torch.int32))
    else:
        if tiles is None:
            _, _, isect_sorted, gids_sorted, tile_bins = bin_and_sort_gaussians(
                num_points, num_intersects, xys, depths, radii, cum, (tiles_x, tiles_y, 1), block_width)
        else:
            # only the requested tiles' lists (same lists, no global key array): see tile_lists
            isect_sorted = None
            gids_sorted, tile_bins = tile_lists(xys, depths, radii, (tiles_x, tiles_y, 1), block_width, tiles)
        out_img, final_ts, final_idx, margin = rasterize_forward(
            xys, conics, colors, opacity, gids_sorted, tile_bins, img_height, img_width,
            block_width, background, tiles=tiles)
        aux = dict(num_intersects=num_intersects, final_idx=final_idx, margin=margin,
                   gaussian_ids_sorted=gids_sorted, isect_ids_sorted=isect_sorted,
                   tile_bins=tile_bins)
    out_alpha = 1.0 - final_ts
    outs = (out_img, out_alpha) if return_alpha else (out_img,)
    if return_aux:
        return outs + (aux,)
    return outs if return_alpha else out_img


# ----------------------------------------------------------------------------
# A.8 wrapper conventions: the caller immediately above the boundary
# ----------------------------------------------------------------------------
def viewmat_from_c2w(camera_to_world: Tensor) -> Tensor:
    """OpenGL camera-to-world (3x4 or 4x4) -> OpenCV world-to-camera 4x4
    (gs_utils.py:32-42: R <- R diag(1,-1,-1); viewmat = [R^T | -R^T T])."""
    rot = camera_to_world[:3, :3] * torch.tensor([1.0, -1.0, -1.0], dtype=camera_to_world.dtype)[None, :]
    trans = camera_to_world[:3, 3]
    vm = torch.eye(4, dtype=camera_to_world.dtype)
    vm[:3, :3] = rot.T
    # -R^T T with every product and sum rounded separately, left to right (arithmetic contract of this
    # file: no matmul -- a BLAS gemv may fuse or reorder); the CUDA kernels repeat exactly this
    for i in range(3):
        vm[i, 3] = -(rot[0, i] * trans[0] + rot[1, i] * trans[1] + rot[2, i] * trans[2])
    return vm


def wrapper_normalise_quats(q: Tensor) -> Tensor:
    """gs_utils.py:46-51, literally: quats = q / norm(q); rows whose result is not unit -- one-sided test
    (norm - 1) < 1e-6, so zero, NaN and infinite quaternions -- are replaced by (0, 0, 0, 1).  The norm
    is the separately rounded left-to-right sum of squares and a correctly rounded sqrt.  Replaced rows
    receive no gradient."""
    def sumsq(t):
        return t[:, 0] * t[:, 0] + t[:, 1] * t[:, 1] + t[:, 2] * t[:, 2] + t[:, 3] * t[:, 3]
    qd = q.detach()
    n = _sqrt(sumsq(qd))
    n1 = _sqrt(sumsq(qd / n[:, None]))
    good = (n1 - 1.0) < 1e-6                                       # False for NaN
    # double-where: the replaced rows go through the differentiable branch as (1,0,0,0), so that their
    # masked-out gradient is an exact zero instead of 0 * inf
    q_safe = torch.where(good[:, None], q, torch.tensor([1.0, 0.0, 0.0, 0.0], dtype=q.dtype).expand_as(q))
    unit = q_safe / _sqrt(sumsq(q_safe))[:, None]
    subst = torch.tensor([0.0, 0.0, 0.0, 1.0], dtype=q.dtype).expand_as(q)
    return torch.where(good[:, None], unit, subst)


def render_view(gs_params: dict, camera_to_world: Tensor, cx, cy, fx, fy, width, height,
                background_color: Tensor, block_width: int = 16, return_aux: bool = False,
                exp=torch.exp, sigmoid=torch.sigmoid, tiles=None):
    """The logic of rasterize_gaussians_to_singleimg (gs_utils.py:29-114) on top of
    this oracle: activations, colour path selection, three gsplat calls, clamps.
    `exp` / `sigmoid`: the activation implementations (gs_utils.py:45,54,65).  The reference runs
    torch.exp / torch.sigmoid ON THE GPU; a parity test that wants the oracle to see the very bits the
    reference's wrapper would hand to gsplat passes differentiable stand-ins that evaluate there
    (tests/helpers.py: gpu_exp, gpu_sigmoid).  `tiles`: composite only these tiles (rasterize_forward)."""
    gp = {k: (v.float() if v.dtype == torch.half else v) for k, v in gs_params.items()}
    viewmat = viewmat_from_c2w(camera_to_world)
    means = gp["means"]
    scales = exp(gp["scales"])
    quats = wrapper_normalise_quats(gp["quats"])
    if "opacities" in gp:
        opacities = sigmoid(gp["opacities"])
    elif "opacities_sigmoid" in gp:
        opacities = gp["opacities_sigmoid"]
    else:
        raise ValueError("No opacities found in gs_params")
    if "features_rest" in gp:
        coeffs = torch.cat([gp["features_dc"].unsqueeze(1), gp["features_rest"]], dim=1)
    else:
        coeffs = gp["features_dc"].unsqueeze(1)
    degree = int(math.sqrt(coeffs.shape[1]) - 1)
    if degree == 0:
        rgbs = sigmoid(coeffs[:, 0, :])
    else:
        dirs = means.detach() - camera_to_world.detach()[:3, 3]
        norm = dirs.norm(dim=-1, keepdim=True)
        # gs_utils.py:72-76: a mean ON the camera centre gets a random unit direction upstream.  Such a
        # Gaussian has depth 0 and is culled by the near plane, so the direction never reaches a pixel; a
        # fixed one keeps the oracle deterministic (and its autograd free of 0 * NaN).
        at_centre = norm == 0
        dirs = torch.where(at_centre, torch.tensor([0.0, 0.0, 1.0], dtype=dirs.dtype).expand_as(dirs),
                           dirs / torch.where(at_centre, torch.ones_like(norm), norm))
        rgbs = torch.clamp(spherical_harmonics(degree, dirs, coeffs) + 0.5, min=0.0)
    h, w = int(height), int(width)
    xys, depths, radii, conics, comp, num_tiles_hit, cov3d = project_gaussians(
        means, scales, 1, quats, viewmat[:3, :].to(means.dtype), float(fx), float(fy), float(cx),
        float(cy), h, w, block_width)
    res = rasterize_gaussians(xys, depths, radii, conics, num_tiles_hit, rgbs, opacities, h, w,
                              block_width, background=background_color.to(means.dtype),
                              return_alpha=True, return_aux=return_aux, tiles=tiles)
    rgb = torch.clamp(res[0], max=1.0)
    alpha = res[1].unsqueeze(-1)
    if return_aux:
        return rgb, alpha, res[2]
    return rgb, alpha


def render_views(gs_params: dict, cameras: dict, block_width: int = 16):
    """rasterize_gaussians_to_multiimgs (gs_utils.py:20-27)."""
    rgbs, alphas = [], []
    kw = {k: cameras[k] for k in ("cx", "cy", "fx", "fy", "width", "height", "background_color")}
    for c2w in cameras["camera_to_worlds"]:
        rgb, alpha = render_view(gs_params, c2w, block_width=block_width, **kw)
        rgbs.append(rgb)
        alphas.append(alpha)
    return rgbs, alphas
