import os, sys, cProfile, pstats
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from moss_b200 import synth
from tests import scenes as S
import moss_b200 as R
dev = torch.device("cuda", 0)
d = S.make_scene("zju")
cams = synth.ring_cameras(23, 512, 512)
names = ("means3D", "shs", "opacities", "cov3D_precomp")
host = {k: d[k].contiguous().pin_memory() for k in names}
rast = [R.GaussianRasterizer(R.GaussianRasterizationSettings(**synth.settings_kwargs(c, d["bg"], d["sh_degree"], d["scale_modifier"], device=dev))) for c in cams]
gc, gd, ga = (t.to(dev) for t in S.upstream_grads(512, 512, depth=False))
def step(i):
    t = {k: host[k].to(dev, non_blocking=True).requires_grad_(True) for k in host}
    means2D = torch.zeros_like(t["means3D"], requires_grad=True)
    color, radii, depth, alpha = rast[i % 23](means3D=t["means3D"], means2D=means2D, shs=t["shs"], colors_precomp=None,
                                              opacities=t["opacities"], scales=None, rotations=None, cov3D_precomp=t["cov3D_precomp"])
    torch.autograd.backward([color, alpha], [gc, ga])
    torch.cuda.synchronize()
for i in range(10): step(i)
pr = cProfile.Profile(); pr.enable()
for i in range(50): step(i)
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(18)
