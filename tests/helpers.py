"""Shared test utilities: the wrapper logic of the reference's render call on a pluggable backend,
comparison helpers and small scene builders."""
from __future__ import annotations

import math

import torch

from oracle import ref_torch as R
from splatformer_b200 import synth

IMG_ATOL = 1e-4        # north_star: rendered RGB / alpha within 1e-4 absolute
GRAD_RTOL = 1e-3       # north_star: per-Gaussian gradients within 1e-3 relative
FRAGILE_MARGIN = 2e-5  # pixels whose alpha>=1/255, T<=1e-4 or 0.999-clamp decision is this close to
                       # its threshold (relative) are set aside: the decision is a discontinuity and
                       # flips with the last bits of exp()


def activated(gs, device=None, dtype=torch.float32):
    """Raw SplatFormer parameters -> the tensors the three gsplat calls receive."""
    means = gs["means"].to(dtype)
    scales = torch.exp(gs["scales"].to(dtype))
    quats = gs["quats"].to(dtype)
    quats = quats / quats.norm(dim=-1, keepdim=True)
    opac = torch.sigmoid(gs["opacities"].to(dtype))
    coeffs = gs["features_dc"].to(dtype).unsqueeze(1)
    if "features_rest" in gs:
        coeffs = torch.cat([coeffs, gs["features_rest"].to(dtype)], dim=1)
    out = dict(means=means, scales=scales, quats=quats, opacities=opac, coeffs=coeffs)
    if device is not None:
        out = {k: v.to(device) for k, v in out.items()}
    return out


def cam_scalars(cams):
    return dict(fx=float(cams["fx"]), fy=float(cams["fy"]), cx=float(cams["cx"]), cy=float(cams["cy"]),
                H=int(cams["height"]), W=int(cams["width"]))


def view_inputs(gs, cams, view=0, device=None, dtype=torch.float32):
    """Everything one per-view pass needs: activated params, viewmat (3x4), unit viewdirs, colours path."""
    a = activated(gs, None, dtype)
    c2w = cams["camera_to_worlds"][view]
    vm = R.viewmat_from_c2w(c2w)[:3, :].contiguous().to(dtype)
    dirs = a["means"] - c2w[:3, 3].to(dtype)
    dirs = dirs / dirs.norm(dim=-1, keepdim=True)
    a.update(viewmat=vm, viewdirs=dirs, background=cams["background_color"].to(dtype))
    if device is not None:
        a = {k: v.to(device) for k, v in a.items()}
    return a


def colors_from(backend, inp, degree):
    if degree == 0:
        return torch.sigmoid(inp["coeffs"][:, 0, :])
    return torch.clamp(backend.spherical_harmonics(degree, inp["viewdirs"], inp["coeffs"]) + 0.5, min=0.0)


def rel_l2(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def assert_grad_close(got, ref, name, rtol=GRAD_RTOL):
    """|got - ref| <= rtol * |ref| + rtol * rms(ref) elementwise, and relative L2 error <= rtol."""
    got, ref = got.double().cpu(), ref.double().cpu()
    assert got.shape == ref.shape, (name, got.shape, ref.shape)
    rms = float(ref.pow(2).mean().sqrt())
    err = (got - ref).abs()
    bound = rtol * ref.abs() + rtol * rms
    bad = err > bound
    assert not bool(bad.any()), (f"{name}: {int(bad.sum())} of {bad.numel()} elements off; worst err "
                                 f"{float(err.max()):.3e}, rms(ref) {rms:.3e}")
    assert rel_l2(got, ref) <= rtol, f"{name}: relative L2 error {rel_l2(got, ref):.3e}"


def small_scene(n=600, deg=2, seed=7, size=64, views=2, scale_mult=0.6):
    gs = synth.shell_scene(n, deg, seed=seed, scale_mult=scale_mult)
    cams = synth.orbit_cameras(views, size, size, background=(0.1, 0.2, 0.3))
    return gs, cams


# ---------------------------------------------------------------------------------------------
# The reference's wrapper evaluates its activations with torch ON THE GPU (gs_utils.py:45,54,65).
# These stand-ins let the CPU oracle see exactly those bits (and differentiate through them), so that
# a parity test of the raw-parameter path compares the kernels with "what gsplat would have been
# handed", not with the CPU libm's last-bit choices.
# ---------------------------------------------------------------------------------------------
class _GpuExp(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        y = torch.exp(x.detach().cuda()).cpu() if x.dtype == torch.float32 else torch.exp(x.detach())
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, g):
        (y,) = ctx.saved_tensors
        return g * y


class _GpuSigmoid(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        y = torch.sigmoid(x.detach().cuda()).cpu() if x.dtype == torch.float32 else torch.sigmoid(x.detach())
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, g):
        (y,) = ctx.saved_tensors
        return g * y * (1.0 - y)


def gpu_exp(x):
    return _GpuExp.apply(x)


def gpu_sigmoid(x):
    return _GpuSigmoid.apply(x)


_REPORT = []


def report(**fields):
    """One line of the parity report (printed, and appended to gpurun_out/parity_report.jsonl when that
    directory exists): what was compared, how many elements, how many were set aside and why."""
    import json
    import os
    line = json.dumps(fields, sort_keys=True)
    print("[parity] " + line)
    _REPORT.append(fields)
    out = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
    if os.path.isdir(out):
        with open(os.path.join(out, "parity_report.jsonl"), "a") as f:
            f.write(line + "\n")


def tile_lists_from_segments(gids_sorted, tile_bins, view, tiles_per_view):
    """The fused path's per-tile lists of one view, concatenated in tile order (what the oracle's
    gaussian_ids_sorted is), and the per-tile lengths."""
    bins = tile_bins[view * tiles_per_view:(view + 1) * tiles_per_view].long().cpu()
    cnt = bins[:, 1] - bins[:, 0]
    total = int(cnt.sum())
    if total == 0:
        return torch.zeros(0, dtype=torch.int32), cnt
    tile_of = torch.repeat_interleave(torch.arange(tiles_per_view), cnt)
    start_out = torch.cumsum(cnt, 0) - cnt
    pos = torch.arange(total) - start_out[tile_of] + bins[tile_of, 0]
    return gids_sorted.cpu()[pos], cnt


def oracle_f64_grads(gs, cams, gt, views, masks=None, extra=None):
    """Gradients of sum_v mean(|rgb_v - gt_v| * mask_v) (+ extra(rgb_v, v)) w.r.t. the raw parameters from the
    FLOAT64 oracle -- the reference for every gradient tolerance (north_star: 1e-3 relative)."""
    raw = {k: v.double().requires_grad_() for k, v in gs.items()}
    kw = {k: cams[k] for k in ("cx", "cy", "fx", "fy", "width", "height", "background_color")}
    loss = 0.0
    for i, v in enumerate(views):
        rgb, _ = R.render_view(raw, cams["camera_to_worlds"][v].double(), **kw)
        m = 1.0 if masks is None else masks[i][..., None].double()
        loss = loss + ((rgb - gt[v].double()).abs() * m).mean()
        if extra is not None:
            loss = loss + extra(rgb, v)
    loss.backward()
    return {k: p.grad for k, p in raw.items()}, float(loss.detach())
