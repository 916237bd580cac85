import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from moss_b200 import synth, _abi
from tests import scenes as S
import moss_b200 as R
dev = torch.device("cuda", 0)
lib = _abi.load()
T = {"fwd_c": 0.0, "bwd_c": 0.0, "n": 0}
class Wrap:
    def __init__(self, lib): self._l = lib
    def __getattr__(self, k):
        f = getattr(self._l, k)
        if k in ("moss_rasterize_forward", "moss_rasterize_backward"):
            key = "fwd_c" if "forward" in k else "bwd_c"
            def g(*a):
                t0 = time.perf_counter(); r = f(*a); T[key] += time.perf_counter() - t0; return r
            return g
        return f
_abi._lib = Wrap(lib)
d = S.make_scene("zju")
cams = synth.ring_cameras(23, 512, 512)
names = ("means3D", "shs", "opacities", "cov3D_precomp")
host = {k: d[k].contiguous().pin_memory() for k in names}
rast = [R.GaussianRasterizer(R.GaussianRasterizationSettings(**synth.settings_kwargs(c, d["bg"], d["sh_degree"], d["scale_modifier"], device=dev))) for c in cams]
gc, gd, ga = (t.to(dev) for t in S.upstream_grads(512, 512, depth=False))
out_host = {}
def step(i, d2h, mode):
    t = {k: host[k].to(dev, non_blocking=True).requires_grad_(True) for k in host}
    means2D = torch.zeros_like(t["means3D"], requires_grad=True)
    color, radii, depth, alpha = rast[i % 23](means3D=t["means3D"], means2D=means2D, shs=t["shs"], colors_precomp=None,
                                              opacities=t["opacities"], scales=None, rotations=None, cov3D_precomp=t["cov3D_precomp"])
    torch.autograd.backward([color, alpha], [gc, ga])
    if d2h:
        res = [color, depth, alpha, radii, means2D.grad] + [t[k].grad for k in host]
        for j, r in enumerate(res):
            if mode == "pinned":
                if j not in out_host:
                    out_host[j] = torch.empty(r.shape, dtype=r.dtype, pin_memory=True)
                out_host[j].copy_(r, non_blocking=True)
            else:
                r.cpu()
    torch.cuda.synchronize()
for label, d2h, mode in (("no d2h", False, ""), ("pinned d2h", True, "pinned"), ("no d2h again", False, ""), ("pageable d2h", True, "cpu")):
    for i in range(10): step(i, d2h, mode)
    T.update(fwd_c=0.0, bwd_c=0.0)
    t0 = time.perf_counter()
    for i in range(50): step(i, d2h, mode)
    dt = (time.perf_counter() - t0) / 50
    print(f"{label:14s} step {dt*1e3:.3f} ms  C fwd call {T['fwd_c']/50*1e3:.3f} ms  C bwd call {T['bwd_c']/50*1e3:.3f} ms", flush=True)
