"""On-disk formats either side of the render path.

  * nerfstudio splatfacto checkpoints `nerfstudio_models/step-*.ckpt` -> the raw Gaussian-parameter
    dict the render path consumes, with the filtering and unit-cube normalisation SplatFormer
    applies when it loads a scene (/root/reference/dataset/GS.py:153-204,
    /root/reference/utils/transform_utils.py:10-42,45-99);
  * Inria-layout binary PLY of a Gaussian set + `cameras.json` / `cfg_args` for the SIBR viewer
    (/root/reference/utils/gs_utils.py:119-261).  The viewer itself is unavailable on this box
    (graphics APIs); the files are what other tools read.  `read_ply` is the inverse, so scenes
    written by the reference (or by Inria 3DGS / nerfstudio exports) can be fed to the renderer.

Host-side file IO only: no kernels, no dependency beyond numpy / torch (the reference uses
`plyfile`; the PLY subset it writes -- one `vertex` element of float32 properties, binary little
endian -- is written and parsed directly here).
"""
from __future__ import annotations

import glob
import json
import math
import os
from argparse import Namespace
from typing import Dict, Iterable, Optional, Tuple

import numpy as np
import torch
from torch import Tensor

C0 = 0.28209479177387814            # SH band-0 constant (gs_utils.py:13)
GAUSS_PREFIX = "_model.gauss_params."
DEFAULT_KEYS = ("means", "scales", "quats", "opacities", "features_dc", "features_rest")


def rgb_to_sh(rgb: Tensor) -> Tensor:
    """gs_utils.py:17 RGB2SH."""
    return (rgb - 0.5) / C0


class MinMaxScaler:
    """Uniform scale + translation that puts a point set into [0,1]^3, centred (the reference's
    MinMaxScaler with preserve_ratio=True, transform_utils.py:45-99).  `scale_` also shifts the
    log-scales of the Gaussians; `transform` maps camera positions into the same frame."""

    def __init__(self):
        self.scale_: Optional[Tensor] = None
        self.trans_: Optional[Tensor] = None

    def fit_transform(self, x: Tensor) -> Tensor:
        lo, hi = x.min(dim=0)[0], x.max(dim=0)[0]
        self.scale_ = torch.min(1.0 / (hi - lo))
        scaled = x * self.scale_
        mid = (scaled.min(dim=0)[0] + scaled.max(dim=0)[0]) / 2
        self.trans_ = 0.5 - mid
        return scaled + self.trans_

    def transform(self, x: Tensor) -> Tensor:
        return x * self.scale_ + self.trans_

    def inverse_transform(self, x: Tensor) -> Tensor:
        return (x - self.trans_) / self.scale_


def remove_outliers(points: Tensor, n_devs: float = 3.0) -> Tuple[Tensor, Tensor]:
    """Keep the points within n_devs standard deviations of the mean on every axis
    (transform_utils.py:10-42, default arguments)."""
    mean, std = points.mean(dim=0), points.std(dim=0)
    mask = torch.all((points >= mean - n_devs * std) & (points <= mean + n_devs * std), dim=1)
    return points[mask], mask


def find_checkpoint(nerfstudio_dir: str) -> str:
    files = sorted(glob.glob(os.path.join(nerfstudio_dir, "nerfstudio_models", "step-*.ckpt")))
    if not files:
        raise FileNotFoundError(f"{nerfstudio_dir} does not have nerfstudio_models/step-*.ckpt")
    return files[-1]                   # the last checkpoint (GS.py:159)


def load_nerfstudio_checkpoint(path: str, keys: Iterable[str] = DEFAULT_KEYS, remove_outlier_ndevs: float = 0.0,
                               max_gs_num: Optional[int] = None, normalize: bool = True
                               ) -> Tuple[Dict[str, Tensor], Optional[MinMaxScaler]]:
    """`path`: a step-*.ckpt file or a nerfstudio output directory.  Returns (gs_params, scaler):
    the `_model.gauss_params.*` tensors named in `keys`, with NaN rows dropped, optional outlier
    removal and truncation, means mapped into [0,1]^3 and log-scales shifted accordingly
    (GS.py:153-204).  normalize=False skips the last step (scaler = None)."""
    ckpt_file = find_checkpoint(path) if os.path.isdir(path) else path
    ckpt = torch.load(ckpt_file, map_location="cpu", weights_only=False)
    if "pipeline" in ckpt and isinstance(ckpt["pipeline"], dict):        # full nerfstudio trainer checkpoint
        ckpt = ckpt["pipeline"]
    found = {k.replace(GAUSS_PREFIX, ""): v for k, v in ckpt.items() if "gauss_params" in k}
    missing = [k for k in keys if k not in found]
    if missing:
        raise KeyError(f"{ckpt_file}: no {GAUSS_PREFIX}{missing[0]} (has {sorted(found)})")
    gs = {k: found[k].detach().float() for k in keys}

    def keep(mask: Tensor) -> None:
        for k in gs:
            gs[k] = gs[k][mask]

    select = torch.ones(gs["means"].shape[0], dtype=torch.bool)
    for k, v in gs.items():
        bad = torch.isnan(v.sum(dim=1)).any(dim=1) if v.dim() == 3 else torch.isnan(v).any(dim=1)
        select &= ~bad
    keep(select)
    if remove_outlier_ndevs > 0:
        keep(remove_outliers(gs["means"], remove_outlier_ndevs)[1])
    if max_gs_num is not None and gs["means"].shape[0] > max_gs_num:
        keep(torch.arange(gs["means"].shape[0]) < max_gs_num)
    scaler = None
    if normalize:
        scaler = MinMaxScaler()
        gs["means"] = scaler.fit_transform(gs["means"])
        if "scales" in gs:
            gs["scales"] = gs["scales"] + torch.log(scaler.scale_)
        valid = torch.all((gs["means"] >= 0) & (gs["means"] <= 1), dim=1)
        if "scales" in gs:
            valid &= ~torch.isinf(gs["scales"]).any(dim=1)
        keep(valid)
    return gs, scaler


# ------------------------------------------------------------------------------------------------
# Inria-layout PLY
# ------------------------------------------------------------------------------------------------
def ply_property_names(num_rest: int) -> list:
    """x y z nx ny nz f_dc_0..2 f_rest_0..(3*num_rest-1) opacity scale_0..2 rot_0..3 (gs_utils.py:199-253)."""
    return (["x", "y", "z", "nx", "ny", "nz"] + [f"f_dc_{i}" for i in range(3)]
            + [f"f_rest_{i}" for i in range(3 * num_rest)] + ["opacity"]
            + [f"scale_{i}" for i in range(3)] + [f"rot_{i}" for i in range(4)])


def export_ply(gs_params: Dict[str, Tensor], filename: str) -> None:
    """export_ply_forviewer (gs_utils.py:156-196): raw parameters as float32 vertex properties.
    features_rest (N,K-1,3) is stored channel-major (transpose(1,2)) as the Inria viewer expects;
    a degree-0 set (no features_rest) stores RGB2SH(sigmoid(features_dc)) as its DC term."""
    d = os.path.dirname(filename)
    if d:
        os.makedirs(d, exist_ok=True)
    with torch.no_grad():
        means = gs_params["means"].detach().float().cpu()
        n = means.shape[0]
        rest = gs_params.get("features_rest")
        if rest is not None and rest.shape[1] != 0:
            f_dc = gs_params["features_dc"].detach().float().cpu().reshape(n, 3)
            f_rest = rest.detach().float().cpu().transpose(1, 2).contiguous().reshape(n, -1)
        else:
            f_dc = rgb_to_sh(torch.sigmoid(gs_params["features_dc"].detach().float().cpu())).reshape(n, 3)
            f_rest = torch.zeros(n, 0)
        cols = [means, torch.zeros(n, 3), f_dc, f_rest,
                gs_params["opacities"].detach().float().cpu().reshape(n, 1),
                gs_params["scales"].detach().float().cpu().reshape(n, 3),
                gs_params["quats"].detach().float().cpu().reshape(n, 4)]
        table = torch.cat(cols, dim=1).numpy().astype("<f4")
    names = ply_property_names(f_rest.shape[1] // 3)
    assert table.shape[1] == len(names)
    header = ["ply", "format binary_little_endian 1.0", f"element vertex {n}"]
    header += [f"property float {name}" for name in names] + ["end_header"]
    with open(filename, "wb") as f:
        f.write(("\n".join(header) + "\n").encode("ascii"))
        f.write(np.ascontiguousarray(table).tobytes())


_PLY_TYPES = {"float": "<f4", "float32": "<f4", "double": "<f8", "float64": "<f8", "uchar": "u1", "uint8": "u1",
              "char": "i1", "int8": "i1", "short": "<i2", "int16": "<i2", "ushort": "<u2", "uint16": "<u2",
              "int": "<i4", "int32": "<i4", "uint": "<u4", "uint32": "<u4"}


def read_ply(filename: str) -> Dict[str, Tensor]:
    """Inverse of export_ply for binary little-endian (or ascii) PLY files with one `vertex` element of
    scalar properties.  Returns means, scales, quats, opacities (N,1), features_dc (N,3) and -- if
    f_rest_* are present -- features_rest (N,K-1,3)."""
    with open(filename, "rb") as f:
        if f.readline().strip() != b"ply":
            raise ValueError(f"{filename} is not a PLY file")
        fmt, count, props, in_vertex = None, None, [], False
        while True:
            line = f.readline()
            if not line:
                raise ValueError(f"{filename}: unterminated PLY header")
            tok = line.decode("ascii").split()
            if not tok or tok[0] == "comment":
                continue
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                in_vertex = tok[1] == "vertex"
                if in_vertex:
                    count = int(tok[2])
                elif count is None:
                    raise ValueError(f"{filename}: elements before `vertex` are not supported")
            elif tok[0] == "property" and in_vertex:
                if tok[1] == "list":
                    raise ValueError(f"{filename}: list properties are not supported")
                props.append((tok[2], _PLY_TYPES[tok[1]]))
            elif tok[0] == "end_header":
                break
        if count is None:
            raise ValueError(f"{filename}: no vertex element")
        if fmt == "binary_little_endian":
            rec = np.frombuffer(f.read(count * np.dtype(props).itemsize), dtype=np.dtype(props), count=count)
            col = {name: rec[name].astype(np.float32) for name, _ in props}
        elif fmt == "ascii":
            arr = np.loadtxt(f, dtype=np.float64, max_rows=count, ndmin=2)
            col = {name: arr[:, i].astype(np.float32) for i, (name, _) in enumerate(props)}
        else:
            raise ValueError(f"{filename}: PLY format {fmt} is not supported")

    def stack(names):
        return torch.from_numpy(np.stack([col[nm] for nm in names], axis=1))

    out = dict(means=stack(["x", "y", "z"]), features_dc=stack([f"f_dc_{i}" for i in range(3)]),
               opacities=stack(["opacity"]), scales=stack([f"scale_{i}" for i in range(3)]),
               quats=stack([f"rot_{i}" for i in range(4)]))
    nrest = sum(1 for name, _ in props if name.startswith("f_rest_"))
    if nrest:
        if nrest % 3:
            raise ValueError(f"{filename}: {nrest} f_rest properties is not a multiple of 3")
        rest = stack([f"f_rest_{i}" for i in range(nrest)])
        out["features_rest"] = rest.reshape(count, 3, nrest // 3).transpose(1, 2).contiguous()
    return out


# ------------------------------------------------------------------------------------------------
# cameras.json / cfg_args for the SIBR viewer
# ------------------------------------------------------------------------------------------------
def focal2fov(focal: float, pixels: float) -> float:
    return 2 * math.atan(pixels / (2 * focal))


def viewer_cameras(cameras: Dict[str, Tensor]) -> list:
    """The cameras.json records of prepare_viewer (gs_utils.py:127-153): per view the OpenCV
    camera-to-world rotation and position (OpenGL c2w with the y and z axes flipped)."""
    out = []
    width, height = int(cameras["width"]), int(cameras["height"])
    fx, fy = float(cameras["fx"]), float(cameras["fy"])
    for i, c2w in enumerate(cameras["camera_to_worlds"]):
        m = np.eye(4)
        m[:3, :4] = c2w.detach().cpu().double().numpy()[:3, :4]
        m[:3, 1:3] *= -1
        # the reference inverts this matrix twice (to w2c, then back through its transposed-R
        # convention): what it stores is the OpenCV camera-to-world matrix itself
        c2w_cv = m
        out.append({"id": i, "img_name": f"img_{i}.png", "width": width, "height": height, "fx": fx, "fy": fy,
                    "FovX": focal2fov(fx, width), "FovY": focal2fov(fy, height),
                    "position": c2w_cv[:3, 3].tolist(), "rotation": [r.tolist() for r in c2w_cv[:3, :3]]})
    return out


def prepare_viewer(cameras: Dict[str, Tensor], dirname: str, sh_degree: int) -> None:
    """Writes `cfg_args` and `cameras.json` next to the exported point cloud (gs_utils.py:119-153)."""
    os.makedirs(dirname, exist_ok=True)
    with open(os.path.join(dirname, "cfg_args"), "w") as f:
        f.write(str(Namespace(source_path="", sh_degree=sh_degree, white_background=False)))
    with open(os.path.join(dirname, "cameras.json"), "w") as f:
        json.dump(viewer_cameras(cameras), f)
