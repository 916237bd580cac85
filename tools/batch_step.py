"""Developer tool: a few fwd+bwd steps of a V-view batch of the ZJU scene on the stream path (for `ncu`)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import moss_b200 as M
from moss_b200 import synth
from tests import scenes as S
dev = torch.device("cuda:0")
V = int(sys.argv[1]) if len(sys.argv) > 1 else 23
d = S.make_scene("zju")
W, H = d["cam"].image_width, d["cam"].image_height
cams = synth.ring_cameras(23, W, H)[:V]
rs = [M.GaussianRasterizationSettings(**synth.settings_kwargs(c, d["bg"], d["sh_degree"], d["scale_modifier"], device=dev)) for c in cams]
t = {k: (d[k].to(dev) if d[k] is not None else torch.empty(0)) for k in ("means3D", "shs", "colors_precomp", "opacities", "scales", "rotations", "cov3D_precomp")}
gs = [S.upstream_grads(W, H, seed=7 + v, depth=False) for v in range(V)]
gc, gd, ga = (torch.stack([g[i] for g in gs]).to(dev) for i in range(3))
for _ in range(int(sys.argv[2]) if len(sys.argv) > 2 else 3):
    c, r, dp, a, st = M.forward_batch_raw(t["means3D"], t["shs"], t["colors_precomp"], t["opacities"], t["scales"], t["rotations"], t["cov3D_precomp"], rs)
    M.backward_batch_raw(st, r, gc, gd, ga)
torch.cuda.synchronize()
print("ok")
