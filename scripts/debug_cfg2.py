import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from oracle import ref_torch as R
from splatformer_b200 import gs_utils, render, synth
import helpers as H
cuda = torch.device("cuda:0")
gs, cams, gt = synth.make_config("objaverse_100k_sh3_512", 0, num_views=32)
view = 5
cams1 = dict(cams); cams1["camera_to_worlds"] = cams["camera_to_worlds"][view:view+1].contiguous()
raw = {k: v.to(cuda) for k, v in gs.items()}
with torch.no_grad():
    rgb, alpha = gs_utils.render_scene(raw, **{k: v.to(cuda) for k, v in cams1.items()})
    render.DEFAULT_TILE_CAPACITY = 1 << 20
    rgb_g, alpha_g = gs_utils.render_scene(raw, **{k: v.to(cuda) for k, v in cams1.items()})
    print("rect vs global-sort path equal:", torch.equal(rgb, rgb_g), float((rgb - rgb_g).abs().max()))
    rgb_ref, a_ref, aux = R.render_view(gs, cams["camera_to_worlds"][view], cams["cx"], cams["cy"], cams["fx"],
                                        cams["fy"], cams["width"], cams["height"], cams["background_color"], return_aux=True)
    gs64 = {k: v.double() for k, v in gs.items()}
    try:
        rgb64, a64, aux64 = R.render_view(gs64, cams["camera_to_worlds"][view].double(), cams["cx"], cams["cy"], cams["fx"],
                                          cams["fy"], cams["width"], cams["height"], cams["background_color"].double(), return_aux=True)
    except Exception as e:
        print("f64 oracle failed", e); rgb64 = None
solid = aux["margin"] > H.FRAGILE_MARGIN
err = (rgb[0].cpu() - rgb_ref).abs().amax(-1)
print("solid frac", float(solid.float().mean()), "err>1e-4:", int(((err > 1e-4) & solid).sum()), "err>5e-5:", int(((err > 5e-5) & solid).sum()),
      "err>2e-5:", int(((err > 2e-5) & solid).sum()))
bad = torch.nonzero((err > 5e-5) & solid)
for y, x in bad[:12].tolist():
    s = f"pix ({y},{x}) err {float(err[y,x]):.2e} margin {float(aux['margin'][y,x]):.2e} gpu {rgb[0,y,x].cpu().tolist()} ref {rgb_ref[y,x].tolist()}"
    if rgb64 is not None:
        s += f" f64 {rgb64[y,x].float().tolist()} margin64 {float(aux64['margin'][y,x]):.2e}"
    print(s)
if rgb64 is not None:
    e_gpu = (rgb[0].cpu().double() - rgb64).abs().amax(-1); e_ref = (rgb_ref.double() - rgb64).abs().amax(-1)
    s64 = aux64["margin"] > H.FRAGILE_MARGIN
    print("vs f64: gpu max", float(e_gpu[s64].max()), "oracle32 max", float(e_ref[s64].max()),
          "gpu>1e-4:", int((e_gpu[s64] > 1e-4).sum()), "oracle32>1e-4:", int((e_ref[s64] > 1e-4).sum()))
