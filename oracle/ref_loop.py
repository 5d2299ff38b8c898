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
