"""On-disk formats either side of the render path: nerfstudio checkpoint loader,
Inria-layout PLY writer / reader, cameras.json for the viewer.  CPU only."""
import json
import math
import os

import numpy as np
import pytest
import torch

from splatformer_b200 import formats, synth


def _fake_checkpoint(tmp_path, n=500, k=16, seed=0):
    gen = torch.Generator().manual_seed(seed)
    gp = dict(means=torch.randn(n, 3, generator=gen) * torch.tensor([2.0, 1.0, 0.5]) + 3.0,
              scales=torch.randn(n, 3, generator=gen) - 4.0, quats=torch.randn(n, 4, generator=gen),
              opacities=torch.randn(n, 1, generator=gen), features_dc=torch.randn(n, 3, generator=gen),
              features_rest=torch.randn(n, k - 1, 3, generator=gen))
    gp["means"][7, 1] = float("nan")
    gp["features_rest"][11, 1, 2] = float("nan")
    gp["means"][20] = torch.tensor([400.0, 3.0, 3.0])             # far outlier
    gp["scales"][30, 0] = float("-inf")
    ckpt = {"_model.gauss_params." + k_: v for k_, v in gp.items()}
    ckpt["_model.camera_optimizer.pose_adjustment"] = torch.zeros(4, 6)
    d = tmp_path / "scene" / "nerfstudio_models"
    d.mkdir(parents=True)
    torch.save(ckpt, d / "step-000009999.ckpt")
    torch.save({"_model.gauss_params.means": torch.zeros(1, 3)}, d / "step-000000100.ckpt")
    return gp, str(tmp_path / "scene")


def test_nerfstudio_checkpoint_loader_filters_and_normalises(tmp_path):
    gp, scene_dir = _fake_checkpoint(tmp_path)
    assert formats.find_checkpoint(scene_dir).endswith("step-000009999.ckpt")          # the last one
    gs, scaler = formats.load_nerfstudio_checkpoint(scene_dir, remove_outlier_ndevs=5.0)
    n = gs["means"].shape[0]
    assert n == 500 - 4                                   # two NaN rows, the outlier, the -inf scale
    assert all(v.shape[0] == n for v in gs.values()) and gs["features_rest"].shape[1:] == (15, 3)
    assert float(gs["means"].min()) >= 0.0 and float(gs["means"].max()) <= 1.0
    lo, hi = gs["means"].min(0)[0], gs["means"].max(0)[0]
    assert torch.allclose((lo + hi) / 2, torch.full((3,), 0.5), atol=1e-6)              # centred
    assert abs(float((hi - lo).max()) - 1.0) < 1e-5                                      # longest axis fills the cube
    # expected values restated from GS.py:163-197 on the surviving rows
    keep = torch.ones(500, dtype=torch.bool)
    keep[[7, 11, 20, 30]] = False
    m = gp["means"][keep]
    scale = torch.min(1.0 / (m.max(0)[0] - m.min(0)[0]))
    assert torch.allclose(scaler.scale_, scale)
    assert torch.allclose(gs["scales"], gp["scales"][keep] + torch.log(scale))
    assert torch.allclose(scaler.inverse_transform(gs["means"]), m, atol=1e-4)
    assert torch.equal(gs["quats"], gp["quats"][keep])
    # truncation and the raw mode
    gs2, sc2 = formats.load_nerfstudio_checkpoint(scene_dir, max_gs_num=100, normalize=False)
    assert sc2 is None and gs2["means"].shape[0] == 100
    with pytest.raises(KeyError):
        formats.load_nerfstudio_checkpoint(os.path.join(scene_dir, "nerfstudio_models", "step-000000100.ckpt"))
    with pytest.raises(FileNotFoundError):
        formats.find_checkpoint(str(tmp_path))


def test_loaded_checkpoint_feeds_the_oracle_renderer(tmp_path):
    from oracle import ref_torch as R
    _, scene_dir = _fake_checkpoint(tmp_path, n=300, k=4, seed=3)
    gs, _ = formats.load_nerfstudio_checkpoint(scene_dir, remove_outlier_ndevs=5.0)
    gs["scales"] = gs["scales"].clamp(max=math.log(0.05))
    cams = synth.orbit_cameras(1, 48, 40)
    rgbs, alphas = R.render_views(gs, cams)
    assert rgbs[0].shape == (40, 48, 3) and torch.isfinite(rgbs[0]).all() and float(alphas[0].max()) > 0


@pytest.mark.parametrize("deg", [0, 1, 3])
def test_ply_round_trip_and_layout(tmp_path, deg):
    gs = synth.shell_scene(257, deg, seed=5)
    path = str(tmp_path / "out" / "point_cloud.ply")
    formats.export_ply(gs, path)
    raw = open(path, "rb").read()
    head, body = raw.split(b"end_header\n", 1)
    lines = head.decode().strip().split("\n")
    k = (deg + 1) ** 2
    names = [ln.split()[2] for ln in lines if ln.startswith("property")]
    assert lines[:3] == ["ply", "format binary_little_endian 1.0", "element vertex 257"]
    assert names == formats.ply_property_names(k - 1) and all(ln.split()[1] == "float" for ln in lines[3:])
    assert names[:9] == ["x", "y", "z", "nx", "ny", "nz", "f_dc_0", "f_dc_1", "f_dc_2"]
    assert names[-8:] == ["opacity", "scale_0", "scale_1", "scale_2", "rot_0", "rot_1", "rot_2", "rot_3"]
    assert len(body) == 257 * len(names) * 4
    table = np.frombuffer(body, dtype="<f4").reshape(257, len(names))
    assert np.array_equal(table[:, 0:3], gs["means"].numpy()) and not table[:, 3:6].any()
    back = formats.read_ply(path)
    for key in ("means", "scales", "quats", "opacities"):
        assert torch.equal(back[key], gs[key]), key
    if deg == 0:
        # degree 0: the DC term holds RGB2SH(sigmoid(features_dc)) (gs_utils.py:178-182)
        assert "features_rest" not in back
        assert torch.allclose(back["features_dc"], (torch.sigmoid(gs["features_dc"]) - 0.5) / formats.C0)
    else:
        assert torch.equal(back["features_dc"], gs["features_dc"])
        assert torch.equal(back["features_rest"], gs["features_rest"])
        # channel-major on disk: f_rest_0..K-2 = red of bases 1..K-1 (the Inria order)
        assert np.array_equal(table[:, 9:9 + k - 1], gs["features_rest"][:, :, 0].numpy())


def test_read_ply_ascii_and_errors(tmp_path):
    p = tmp_path / "a.ply"
    names = formats.ply_property_names(0)
    rows = np.arange(2 * len(names), dtype=np.float32).reshape(2, len(names))
    p.write_text("ply\nformat ascii 1.0\ncomment made by hand\nelement vertex 2\n"
                 + "".join(f"property float {n}\n" for n in names) + "end_header\n"
                 + "\n".join(" ".join(str(float(x)) for x in r) for r in rows) + "\n")
    back = formats.read_ply(str(p))
    assert torch.equal(back["means"], torch.from_numpy(rows[:, 0:3]))
    assert torch.equal(back["quats"], torch.from_numpy(rows[:, -4:]))
    (tmp_path / "bad.ply").write_text("not a ply\n")
    with pytest.raises(ValueError):
        formats.read_ply(str(tmp_path / "bad.ply"))


def test_viewer_cameras_match_the_reference_arithmetic(tmp_path):
    cams = synth.orbit_cameras(3, 64, 48)
    formats.prepare_viewer(cams, str(tmp_path / "v"), sh_degree=3)
    recs = json.load(open(tmp_path / "v" / "cameras.json"))
    assert open(tmp_path / "v" / "cfg_args").read() == "Namespace(source_path='', sh_degree=3, white_background=False)"
    assert [r["id"] for r in recs] == [0, 1, 2] and recs[1]["img_name"] == "img_1.png"
    for i, r in enumerate(recs):
        # gs_utils.py:137-150 restated literally
        c2w = np.eye(4)
        c2w[:3, :4] = cams["camera_to_worlds"][i].numpy()[:3, :4]
        c2w[:3, 1:3] *= -1
        w2c = np.linalg.inv(c2w)
        rot_t = np.transpose(w2c[:3, :3])
        rt = np.zeros((4, 4)); rt[:3, :3] = rot_t.transpose(); rt[:3, 3] = w2c[:3, 3]; rt[3, 3] = 1.0
        back = np.linalg.inv(rt)
        assert np.allclose(r["position"], back[:3, 3], atol=1e-12)
        assert np.allclose(r["rotation"], back[:3, :3], atol=1e-12)
        assert r["width"] == 64 and r["height"] == 48
        assert math.isclose(r["FovX"], 2 * math.atan(64 / (2 * float(cams["fx"]))))
