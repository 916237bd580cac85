import os, sys, gc
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from moss_b200 import synth
from tests import scenes as S
impl = sys.argv[1]
if impl == "ours":
    import moss_b200 as R
else:
    sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref")); import diff_gaussian_rasterization as R
dev = torch.device("cuda", 0)
d = S.make_scene("zju")
cams = synth.ring_cameras(23, 512, 512)
names = ("means3D", "shs", "opacities", "cov3D_precomp")
host = {k: d[k].contiguous().pin_memory() for k in names}
rast = [R.GaussianRasterizer(R.GaussianRasterizationSettings(**synth.settings_kwargs(c, d["bg"], d["sh_degree"], d["scale_modifier"], device=dev))) for c in cams]
gc_, gd, ga = (t.to(dev) for t in S.upstream_grads(512, 512, depth=False))
def step(i):
    t = {k: host[k].to(dev, non_blocking=True).requires_grad_(True) for k in host}
    means2D = torch.zeros_like(t["means3D"], requires_grad=True)
    color, radii, depth, alpha = rast[i % 23](means3D=t["means3D"], means2D=means2D, shs=t["shs"], colors_precomp=None,
                                              opacities=t["opacities"], scales=None, rotations=None, cov3D_precomp=t["cov3D_precomp"])
    torch.autograd.backward([color, alpha], [gc_, ga])
    torch.cuda.synchronize()
gc.collect(); gc.disable()
for i in range(5): step(i)
base = torch.cuda.memory_allocated()
for i in range(20): step(i)
grown = torch.cuda.memory_allocated() - base
n = gc.collect()
after = torch.cuda.memory_allocated() - base
st = torch.cuda.memory_stats()
print(impl, f"growth over 20 steps without GC: {grown/1e6:.1f} MB; after gc.collect() ({n} objects): {after/1e6:.1f} MB; device mallocs {st['num_device_alloc']}")
if grown > 1e6:
    gc.set_debug(gc.DEBUG_SAVEALL)
    step(0); gc.collect()
    kinds = {}
    for o in gc.garbage: kinds[type(o).__name__] = kinds.get(type(o).__name__, 0) + 1
    print("garbage types:", sorted(kinds.items(), key=lambda kv: -kv[1])[:12])
