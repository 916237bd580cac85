"""Developer tool: a few launches of the kernels around the rasterizer (nearest vertex, skinning, densification
statistics, AdamW) at ZJU size, for `ncu` captures."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import moss_b200 as M
from moss_b200 import skinning as K, synth
from moss_b200.optim import FusedAdamW

dev = torch.device("cuda:0")
P = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
body = synth.body()
g = synth.gaussians(P, body, seed=synth.SEED + 5)
xyz, verts = g.xyz.to(dev), body.verts.to(dev)
radii = (torch.rand(P) * 20 - 4).clamp_min(0).to(torch.int32).to(dev)
vgrad = torch.randn(P, 3, device=dev)
acc, den, mx = torch.zeros(P, 1, device=dev), torch.zeros(P, 1, device=dev), torch.zeros(P, device=dev)
shapes = {"xyz": 3, "f_dc": 3, "f_rest": 45, "opacity": 1, "scaling": 3, "rotation": 4}
groups = [{"params": [torch.nn.Parameter(torch.randn(P, c, device=dev))], "lr": 1e-3, "name": n} for n, c in shapes.items()]
opt = FusedAdamW(groups, lr=0.0, eps=1e-15)
for it in range(4):
    ids = K.nearest_vertex(xyz, verts)
    M.add_densification_stats(radii, vgrad, acc, den, mx)
    for grp in groups:
        grp["params"][0].grad = torch.randn_like(grp["params"][0])
    opt.step()
torch.cuda.synchronize()
print("ok", int(ids.sum()))
