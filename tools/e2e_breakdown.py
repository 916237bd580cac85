"""Developer tool: where does the ZJU e2e step (host buffers in, image + gradients out) spend its time?
Times each leg alone, synchronised: H2D of the inputs, the public-API fwd+bwd on resident inputs, D2H of the outputs."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import moss_b200 as M
from moss_b200 import synth
from tests import scenes as S

dev = torch.device("cuda:0")
d = S.make_scene("zju")
cam = d["cam"]
rs = M.GaussianRasterizationSettings(**synth.settings_kwargs(cam, d["bg"], d["sh_degree"], d["scale_modifier"], device=dev))
rast = M.GaussianRasterizer(rs)
names = [k for k in ("means3D", "shs", "colors_precomp", "opacities", "scales", "rotations", "cov3D_precomp") if d[k] is not None]
host = {k: d[k].contiguous().pin_memory() for k in names}
gc, gd, ga = (x.to(dev) for x in S.upstream_grads(cam.image_width, cam.image_height, depth=False))
devt = {k: host[k].to(dev) for k in names}
out_host = {}


def h2d():
    return {k: host[k].to(dev, non_blocking=True).requires_grad_(True) for k in names}


def compute(t):
    m2 = torch.zeros_like(t["means3D"], requires_grad=True)
    c, r, dp, a = rast(means3D=t["means3D"], means2D=m2, shs=t.get("shs"), colors_precomp=t.get("colors_precomp"),
                       opacities=t["opacities"], scales=t.get("scales"), rotations=t.get("rotations"), cov3D_precomp=t.get("cov3D_precomp"))
    torch.autograd.backward([c, a], [gc, ga])
    return [c, dp, a, r, m2.grad] + [t[k].grad for k in names]


def d2h(res, base=0):
    for j, r in enumerate(res, base):
        if j not in out_host:
            out_host[j] = torch.empty(r.shape, dtype=r.dtype, pin_memory=True)
        out_host[j].copy_(r.detach(), non_blocking=True)


def timeit(label, fn, n=200):
    for _ in range(10):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
        torch.cuda.current_stream().synchronize()
    dt = (time.perf_counter() - t0) / n * 1e6
    print(f"{label:34s} {dt:8.1f} us/step (wall, synchronised every step)")


res0 = compute({k: devt[k].detach().requires_grad_(True) for k in names})
timeit("h2d only", h2d)
timeit("fwd+bwd (resident inputs)", lambda: compute({k: devt[k].detach().requires_grad_(True) for k in names}))
timeit("d2h only", lambda: d2h(res0))
timeit("d2h grads only", lambda: d2h(res0[4:], 4))
def raw():
    c, r, dp, a, st = M.forward_raw(devt["means3D"], devt.get("shs", E), devt.get("colors_precomp", E), devt["opacities"],
                                    devt.get("scales", E), devt.get("rotations", E), devt.get("cov3D_precomp", E), rs)
    M.backward_raw(st, r, gc, gd, ga)


def fwd_only():
    t = {k: devt[k].detach().requires_grad_(True) for k in names}
    m2 = torch.zeros_like(t["means3D"], requires_grad=True)
    return rast(means3D=t["means3D"], means2D=m2, shs=t.get("shs"), colors_precomp=t.get("colors_precomp"),
                opacities=t["opacities"], scales=t.get("scales"), rotations=t.get("rotations"), cov3D_precomp=t.get("cov3D_precomp"))


E = torch.empty(0, device=dev)
timeit("forward_raw + backward_raw", raw)
timeit("public forward only", fwd_only)
timeit("requires_grad_ x4 + zeros_like", lambda: ({k: devt[k].detach().requires_grad_(True) for k in names}, torch.zeros_like(devt["means3D"], requires_grad=True)))
timeit("h2d + fwd+bwd", lambda: compute(h2d()))
timeit("all three", lambda: d2h(compute(h2d())))
for k in names:
    timeit(f"h2d {k} ({host[k].numel() * 4 >> 10} KiB)", lambda k=k: host[k].to(dev, non_blocking=True))
