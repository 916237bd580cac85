"""Developer tool: a few fwd+bwd steps of one scene on the stream path (for `ncu` launch lists)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import moss_b200 as M
from moss_b200 import synth
from tests import scenes as S
dev = torch.device("cuda:0")
name = sys.argv[1] if len(sys.argv) > 1 else "tiny"
d = S.make_scene(name)
cam = d["cam"]
rs = M.GaussianRasterizationSettings(**synth.settings_kwargs(cam, d["bg"], d["sh_degree"], d["scale_modifier"], device=dev))
t = {k: (d[k].to(dev) if d[k] is not None else torch.empty(0)) for k in ("means3D", "shs", "colors_precomp", "opacities", "scales", "rotations", "cov3D_precomp")}
gc, gd, ga = (x.to(dev) for x in S.upstream_grads(cam.image_width, cam.image_height, depth=False))
for _ in range(int(sys.argv[2]) if len(sys.argv) > 2 else 6):
    c, r, dp, a, st = M.forward_raw(t["means3D"], t["shs"], t["colors_precomp"], t["opacities"], t["scales"], t["rotations"], t["cov3D_precomp"], rs)
    M.backward_raw(st, r, gc, gd, ga)
torch.cuda.synchronize()
print("ok")
