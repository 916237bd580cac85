"""Host-side phase breakdown of one e2e step (H2D, forward, backward, D2H) for ours and the reference arm."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from moss_b200 import synth
from tests import scenes as S

def run(impl, n=100):
    dev = torch.device("cuda", 0)
    if impl == "ours":
        import moss_b200 as R
    else:
        sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))
        import diff_gaussian_rasterization as R
    d = S.make_scene("zju")
    cams = synth.ring_cameras(23, 512, 512)
    names = ("means3D", "shs", "colors_precomp", "opacities", "scales", "rotations", "cov3D_precomp")
    host = {k: d[k].contiguous().pin_memory() for k in names if d[k] is not None}
    rast = [R.GaussianRasterizer(R.GaussianRasterizationSettings(**synth.settings_kwargs(c, d["bg"], d["sh_degree"], d["scale_modifier"], device=dev))) for c in cams]
    gc, gd, ga = (t.to(dev) for t in S.upstream_grads(512, 512, depth=False))
    out_host = {}
    acc = dict(h2d=0.0, fwd_host=0.0, fwd=0.0, bwd_host=0.0, bwd=0.0, d2h=0.0, total=0.0)
    sync = lambda: torch.cuda.synchronize(dev)
    for i in range(n + 10):
        if i == 10:
            acc = {k: 0.0 for k in acc}
        sync(); t0 = time.perf_counter()
        t = {k: host[k].to(dev, non_blocking=True).requires_grad_(True) for k in host}
        means2D = torch.zeros_like(t["means3D"], requires_grad=True)
        sync(); t1 = time.perf_counter()
        color, radii, depth, alpha = rast[i % 23](means3D=t["means3D"], means2D=means2D, shs=t.get("shs"), colors_precomp=None,
                                                  opacities=t["opacities"], scales=None, rotations=None, cov3D_precomp=t.get("cov3D_precomp"))
        t2h = time.perf_counter(); sync(); t2 = time.perf_counter()
        torch.autograd.backward([color, alpha], [gc, ga])
        t3h = time.perf_counter(); sync(); t3 = time.perf_counter()
        res = [color, depth, alpha, radii, means2D.grad] + [t[k].grad for k in host]
        for j, r in enumerate(res):
            if j not in out_host:
                out_host[j] = torch.empty(r.shape, dtype=r.dtype, pin_memory=True)
            out_host[j].copy_(r, non_blocking=True)
        sync(); t4 = time.perf_counter()
        for k, v in (("h2d", t1 - t0), ("fwd_host", t2h - t1), ("fwd", t2 - t1), ("bwd_host", t3h - t2), ("bwd", t3 - t2), ("d2h", t4 - t3), ("total", t4 - t0)):
            acc[k] += v
    print(impl, {k: round(v / n * 1e3, 3) for k, v in acc.items()}, "ms", flush=True)
    print("  grad strides", [(tuple(r.shape), r.is_contiguous()) for r in res[4:]])

if __name__ == "__main__":
    run(sys.argv[1])
