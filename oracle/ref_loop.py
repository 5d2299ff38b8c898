"""ORACLE (test infrastructure, not product code) -- parity unpinned.

Literal per-pixel loops for A.5 (rasterise forward) and A.6 (rasterise backward)
of SURVEY.md Appendix A, in float64 Python scalars.  It exists to cross-check
the vectorised oracle (`oracle/ref_torch.py`: masked cumprod + autograd) on tiny
cases (<= a few hundred Gaussians, <= 48x48 pixels): the forward here is the
sequential walk with `continue` / `break`, the backward is the hand-derived
reverse walk, so an error in the cumprod mask algebra or in the hand-derived
VJP shows up as a disagreement between the two.

Reference boundary: gsplat.rasterize_gaussians as called at
`/root/reference/utils/gs_utils.py:96-109` (gsplat v0.1.11 is not on disk; see
`oracle/ref_torch.py` header).
"""
from __future__ import annotations

import math

import numpy as np

ALPHA_MIN = 1.0 / 255.0
ALPHA_MAX = 0.999
T_MIN = 1e-4


def rasterize_loop(xys, conics, colors, opacity, gaussian_ids_sorted, tile_bins, img_height,
                   img_width, block_width, background, v_out_img=None, v_out_alpha=None):
    """Forward (and, if v_out_img is given, backward) over numpy float64 arrays.

    Returns dict(out_img, out_alpha, final_Ts, final_idx[, v_xy, v_conic, v_colors, v_opacity]).
    """
    xys = np.asarray(xys, dtype=np.float64)
    conics = np.asarray(conics, dtype=np.float64)
    colors = np.asarray(colors, dtype=np.float64)
    opacity = np.asarray(opacity, dtype=np.float64).reshape(-1)
    background = np.asarray(background, dtype=np.float64)
    ids = np.asarray(gaussian_ids_sorted, dtype=np.int64)
    bins = np.asarray(tile_bins, dtype=np.int64)
    tiles_x = (img_width + block_width - 1) // block_width
    n = xys.shape[0]

    out_img = np.zeros((img_height, img_width, 3))
    final_ts = np.ones((img_height, img_width))
    final_idx = np.zeros((img_height, img_width), dtype=np.int64)
    do_bwd = v_out_img is not None
    if do_bwd:
        v_out_img = np.asarray(v_out_img, dtype=np.float64)
        v_out_alpha = (np.zeros((img_height, img_width)) if v_out_alpha is None
                       else np.asarray(v_out_alpha, dtype=np.float64))
        v_xy = np.zeros((n, 2))
        v_conic = np.zeros((n, 3))
        v_colors = np.zeros((n, 3))
        v_opacity = np.zeros(n)

    for i in range(img_height):
        for j in range(img_width):
            tile = (i // block_width) * tiles_x + (j // block_width)
            lo, hi = int(bins[tile, 0]), int(bins[tile, 1])
            px, py = j + 0.5, i + 0.5
            t = 1.0
            cur = 0
            pix = np.zeros(3)
            for idx in range(lo, hi):
                g = ids[idx]
                dx, dy = xys[g, 0] - px, xys[g, 1] - py
                sigma = 0.5 * (conics[g, 0] * dx * dx + conics[g, 2] * dy * dy) + conics[g, 1] * dx * dy
                alpha = min(ALPHA_MAX, opacity[g] * math.exp(-sigma))
                if sigma < 0 or alpha < ALPHA_MIN:
                    continue
                next_t = t * (1.0 - alpha)
                if next_t <= T_MIN:
                    break
                pix += colors[g] * (alpha * t)
                t = next_t
                cur = idx
            final_ts[i, j] = t
            final_idx[i, j] = cur
            out_img[i, j] = pix + t * background

            if not do_bwd:
                continue
            # ---- A.6: walk back from the last composited Gaussian ----------
            v_out = v_out_img[i, j]
            v_oa = v_out_alpha[i, j]
            t_final = t
            tt = t_final
            buf = np.zeros(3)
            for idx in range(cur, lo - 1, -1):
                if hi <= lo:
                    break
                g = ids[idx]
                dx, dy = xys[g, 0] - px, xys[g, 1] - py
                sigma = 0.5 * (conics[g, 0] * dx * dx + conics[g, 2] * dy * dy) + conics[g, 1] * dx * dy
                vis = math.exp(-sigma)
                alpha_raw = opacity[g] * vis
                alpha = min(ALPHA_MAX, alpha_raw)
                if sigma < 0 or alpha < ALPHA_MIN:
                    continue
                ra = 1.0 / (1.0 - alpha)
                tt *= ra                       # T before this Gaussian
                fac = alpha * tt
                v_colors[g] += fac * v_out
                v_alpha = float(np.dot(colors[g] * tt - buf * ra, v_out))
                # out_alpha = 1 - T_final  => d out_alpha / d alpha = +T_final * ra
                v_alpha += t_final * ra * v_oa
                # background term: out_img += T_final * background
                v_alpha += -t_final * ra * float(np.dot(background, v_out))
                buf += colors[g] * fac
                if alpha_raw > ALPHA_MAX:
                    continue                   # clamp active: no gradient to sigma / opacity
                v_sigma = -opacity[g] * vis * v_alpha
                v_conic[g, 0] += 0.5 * v_sigma * dx * dx
                v_conic[g, 1] += v_sigma * dx * dy
                v_conic[g, 2] += 0.5 * v_sigma * dy * dy
                v_xy[g, 0] += v_sigma * (conics[g, 0] * dx + conics[g, 1] * dy)
                v_xy[g, 1] += v_sigma * (conics[g, 1] * dx + conics[g, 2] * dy)
                v_opacity[g] += vis * v_alpha

    res = dict(out_img=out_img, out_alpha=1.0 - final_ts, final_Ts=final_ts, final_idx=final_idx)
    if do_bwd:
        res.update(v_xy=v_xy, v_conic=v_conic, v_colors=v_colors, v_opacity=v_opacity)
    return res


# ----------------------------------------------------------------------------------------------
# A.2 / A.3 / A.4 once more, as a literal per-Gaussian scalar loop in binary32 (numpy float32 scalars): a
# second, independently structured statement of the projection and of the tiling -- 3x3 matrices and explicit
# matrix products the way a glm-based kernel is written, where oracle/ref_torch.py is an unrolled list of
# vector expressions.  Every constant of A.2 is load-bearing for bit-exact tiling (0.3 blur, 1.3 x tan(fov/2)
# clamp, the 0.1 floor of the discriminant, ceil(3 sqrt(lambda)), the 1e-6 in the pixel centre, truncation toward
# zero in the tile bbox); tests/test_oracle_crosscheck.py demands that the two statements agree BIT FOR BIT.
# ----------------------------------------------------------------------------------------------
F = np.float32


def _dot3(a, b):
    """a[0]*b[0] + a[1]*b[1] + a[2]*b[2], every product and sum rounded to binary32, left to right."""
    return F(F(F(a[0] * b[0]) + F(a[1] * b[1])) + F(a[2] * b[2]))


def _trunc_clamp(v, hi):
    if not np.isfinite(v):
        iv = 0 if np.isnan(v) else (2000000000 if v > 0 else -2000000000)
    else:
        iv = int(np.clip(v, -2.0e9, 2.0e9))          # int() truncates toward zero
    return min(max(iv, 0), hi)


def project_loop(means3d, scales, glob_scale, quats, viewmat, fx, fy, cx, cy, img_height, img_width, block_width,
                 clip_thresh=0.01):
    """gsplat.project_gaussians (call site /root/reference/utils/gs_utils.py:82-95), one Gaussian at a time."""
    means3d = np.asarray(means3d, dtype=np.float32)
    scales = np.asarray(scales, dtype=np.float32)
    quats = np.asarray(quats, dtype=np.float32)
    vm = np.asarray(viewmat, dtype=np.float32)
    n = means3d.shape[0]
    fx, fy, cx, cy, clip, glob = F(fx), F(fy), F(cx), F(cy), F(clip_thresh), F(glob_scale)
    W = vm[:3, :3]
    p = vm[:3, 3]
    xys = np.zeros((n, 2), np.float32); depths = np.zeros(n, np.float32); radii = np.zeros(n, np.int32)
    conics = np.zeros((n, 3), np.float32); comp = np.zeros(n, np.float32); tiles = np.zeros(n, np.int32)
    cov3d = np.zeros((n, 6), np.float32)
    tiles_x = (img_width + block_width - 1) // block_width
    tiles_y = (img_height + block_width - 1) // block_width
    lim_x = F(F(1.3) * F(F(F(0.5) * F(img_width)) / fx))
    lim_y = F(F(1.3) * F(F(F(0.5) * F(img_height)) / fy))
    with np.errstate(all="ignore"):
        for g in range(n):
            m = means3d[g]
            t = [F(_dot3(W[i], m) + p[i]) for i in range(3)]                   # camera-space mean
            if not (t[2] > clip):
                continue
            q = quats[g]
            qn = np.sqrt(F(F(F(F(q[0] * q[0]) + F(q[1] * q[1])) + F(q[2] * q[2])) + F(q[3] * q[3])))
            w_, x_, y_, z_ = (F(q[k] / qn) for k in range(4))
            two, one = F(2.0), F(1.0)
            R = [[F(one - F(two * F(F(y_ * y_) + F(z_ * z_)))), F(two * F(F(x_ * y_) - F(w_ * z_))), F(two * F(F(x_ * z_) + F(w_ * y_)))],
                 [F(two * F(F(x_ * y_) + F(w_ * z_))), F(one - F(two * F(F(x_ * x_) + F(z_ * z_)))), F(two * F(F(y_ * z_) - F(w_ * x_)))],
                 [F(two * F(F(x_ * z_) - F(w_ * y_))), F(two * F(F(y_ * z_) + F(w_ * x_))), F(one - F(two * F(F(x_ * x_) + F(y_ * y_))))]]
            s = [F(glob * scales[g][j]) for j in range(3)]
            M = [[F(R[i][j] * s[j]) for j in range(3)] for i in range(3)]        # R S
            Sig = [[_dot3(M[i], M[j]) for j in range(3)] for i in range(3)]      # (R S)(R S)^T
            cov3d[g] = (Sig[0][0], Sig[0][1], Sig[0][2], Sig[1][1], Sig[1][2], Sig[2][2])
            rz = F(one / t[2]); rz2 = F(rz * rz)
            txc = F(t[2] * min(lim_x, max(F(-lim_x), F(t[0] * rz))))
            tyc = F(t[2] * min(lim_y, max(F(-lim_y), F(t[1] * rz))))
            J = [[F(fx * rz), F(0.0), F(F(-F(fx * txc)) * rz2)],
                 [F(0.0), F(fy * rz), F(F(-F(fy * tyc)) * rz2)]]
            # T = J W (the zero entries of J drop out of the sums, as in the unrolled statement)
            T = [[F(F(J[0][0] * W[0][j]) + F(J[0][2] * W[2][j])) for j in range(3)],
                 [F(F(J[1][1] * W[1][j]) + F(J[1][2] * W[2][j])) for j in range(3)]]
            A = [[_dot3(T[r], [Sig[0][j], Sig[1][j], Sig[2][j]]) for j in range(3)] for r in range(2)]   # T Sigma
            c00, c01, c11 = _dot3(A[0], T[0]), _dot3(A[0], T[1]), _dot3(A[1], T[1])
            det_orig = F(F(c00 * c11) - F(c01 * c01))
            a, b, c = F(c00 + F(0.3)), c01, F(c11 + F(0.3))
            det = F(F(a * c) - F(b * b))
            if det == 0 or not np.isfinite(det):
                continue
            ratio = F(det_orig / det)
            cmp_ = np.sqrt(ratio) if ratio > 0 else F(0.0)
            inv_det = F(one / det)
            conic = (F(c * inv_det), F(F(-b) * inv_det), F(a * inv_det))
            bmid = F(F(0.5) * F(a + c))
            disc = np.sqrt(max(F(0.1), F(F(bmid * bmid) - det)))
            radius = np.ceil(F(F(3.0) * np.sqrt(max(F(bmid + disc), F(bmid - disc)))))
            rw = F(one / F(t[2] + F(1e-6)))
            px = F(F(fx * F(t[0] * rw)) + cx)
            py = F(F(fy * F(t[1] * rw)) + cy)
            if not (np.isfinite(radius) and radius <= 1073741824.0 and np.isfinite(px) and np.isfinite(py)):
                continue
            bw = F(block_width)
            tcx, tcy, tr = F(px / bw), F(py / bw), F(radius / bw)
            min_x, max_x = _trunc_clamp(F(tcx - tr), tiles_x), _trunc_clamp(F(F(tcx + tr) + one), tiles_x)
            min_y, max_y = _trunc_clamp(F(tcy - tr), tiles_y), _trunc_clamp(F(F(tcy + tr) + one), tiles_y)
            area = (max_x - min_x) * (max_y - min_y)
            if area <= 0:
                continue
            xys[g] = (px, py); depths[g] = t[2]; radii[g] = int(radius); conics[g] = conic; comp[g] = cmp_
            tiles[g] = area
    return xys, depths, radii, conics, comp, tiles, cov3d


def bin_and_sort_loop(xys, depths, radii, tile_bounds, block_width):
    """A.4 as loops: emit (tile << 32 | depth bits, id) per covered tile in Gaussian order, stable sort by key,
    bin edges by a scan over the sorted keys.  Returns (keys_sorted, ids_sorted, tile_bins)."""
    tiles_x, tiles_y = int(tile_bounds[0]), int(tile_bounds[1])
    pairs = []
    bw = F(block_width)
    for g in range(len(radii)):
        if radii[g] <= 0:
            continue
        tcx, tcy, tr = F(F(xys[g][0]) / bw), F(F(xys[g][1]) / bw), F(F(radii[g]) / bw)
        min_x, max_x = _trunc_clamp(F(tcx - tr), tiles_x), _trunc_clamp(F(F(tcx + tr) + F(1.0)), tiles_x)
        min_y, max_y = _trunc_clamp(F(tcy - tr), tiles_y), _trunc_clamp(F(F(tcy + tr) + F(1.0)), tiles_y)
        dbits = int(np.float32(depths[g]).view(np.int32)) & 0xFFFFFFFF
        for i in range(min_y, max_y):
            for j in range(min_x, max_x):
                pairs.append((((i * tiles_x + j) << 32) | dbits, g))
    pairs.sort(key=lambda kv: kv[0])                      # Python's sort is stable
    keys = np.array([k for k, _ in pairs], dtype=np.int64)
    ids = np.array([g for _, g in pairs], dtype=np.int32)
    bins = np.zeros((tiles_x * tiles_y, 2), dtype=np.int32)
    for idx in range(len(pairs)):
        cur = int(keys[idx] >> 32)
        if idx == 0:
            bins[cur][0] = 0
        elif int(keys[idx - 1] >> 32) != cur:
            bins[int(keys[idx - 1] >> 32)][1] = idx
            bins[cur][0] = idx
        if idx == len(pairs) - 1:
            bins[cur][1] = len(pairs)
    return keys, ids, bins
