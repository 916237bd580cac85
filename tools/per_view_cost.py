"""Developer tool: fwd+bwd device time and instance count of each of the 23 ring cameras (view-cost spread, which is
what a data-parallel step waits for)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import moss_b200 as M
from moss_b200 import synth
from tests import scenes as S
dev = torch.device("cuda:0")
d = S.make_scene(sys.argv[1] if len(sys.argv) > 1 else "zju")
W, H = d["cam"].image_width, d["cam"].image_height
cams = synth.ring_cameras(23, W, H)
t = {k: (d[k].to(dev) if d[k] is not None else torch.empty(0)) for k in ("means3D", "shs", "colors_precomp", "opacities", "scales", "rotations", "cov3D_precomp")}
gc, gd, ga = (x.to(dev) for x in S.upstream_grads(W, H, depth=False))
flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=dev)
res = []
for ci, cam in enumerate(cams):
    rs = M.GaussianRasterizationSettings(**synth.settings_kwargs(cam, d["bg"], d["sh_degree"], d["scale_modifier"], device=dev))
    def step():
        c, r, dp, a, st = M.forward_raw(t["means3D"], t["shs"], t["colors_precomp"], t["opacities"], t["scales"], t["rotations"], t["cov3D_precomp"], rs)
        M.backward_raw(st, r, gc, gd, ga)
        return st
    for _ in range(5): st = step()
    tot = 0.0
    for _ in range(30):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); st = step(); e1.record(); torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    rng = M.export_state(st, "ranges")
    res.append((ci, tot / 30 * 1e3, st.num_rendered, int((rng[:, 1] - rng[:, 0]).max())))
for r in res: print("cam %2d: %6.1f us  R=%6d  deepest tile=%5d" % r)
ts = [r[1] for r in res]
print("mean %.1f  min %.1f  max %.1f" % (sum(ts) / len(ts), min(ts), max(ts)))
