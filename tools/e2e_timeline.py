"""Developer tool: GPU and host timeline of the ZJU e2e step (the loop bench.py's `e2e` times): CUDA events at the
milestones of one step on the streams that carry them, host clocks at the points where the Python call path hands work
over.  Prints mean offsets (us) from the start of the step."""
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
out_host = {}
side = torch.cuda.Stream(dev)
main = torch.cuda.current_stream(dev)
flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=dev)
EV = ["start", "h2d_done", "fwd_done", "img_d2h_done", "bwd_done", "grad_d2h_done"]
HT = ["h2d_issued", "fwd_returned", "img_copies_issued", "bwd_returned", "grad_copies_issued", "synced"]


def to_host(j, r):
    if j not in out_host:
        out_host[j] = torch.empty(r.shape, dtype=r.dtype, pin_memory=True)
    out_host[j].copy_(r.detach(), non_blocking=True)


def step():
    ev = {k: torch.cuda.Event(enable_timing=True) for k in EV}
    ht = {}
    t0 = time.perf_counter()
    ev["start"].record(main)
    t = {k: host[k].to(dev, non_blocking=True).requires_grad_(True) for k in names}
    ev["h2d_done"].record(main); ht["h2d_issued"] = time.perf_counter()
    m2 = torch.zeros_like(t["means3D"], requires_grad=True)
    c, r, dp, a = rast(means3D=t["means3D"], means2D=m2, shs=t.get("shs"), colors_precomp=t.get("colors_precomp"),
                       opacities=t["opacities"], scales=t.get("scales"), rotations=t.get("rotations"), cov3D_precomp=t.get("cov3D_precomp"))
    ev["fwd_done"].record(main); ht["fwd_returned"] = time.perf_counter()
    side.wait_stream(main)
    with torch.cuda.stream(side):
        for j, x in enumerate((c, dp, a, r)):
            to_host(j, x)
        ev["img_d2h_done"].record(side)
    ht["img_copies_issued"] = time.perf_counter()
    torch.autograd.backward([c, a], [gc, ga])
    ev["bwd_done"].record(main); ht["bwd_returned"] = time.perf_counter()
    for j, g in enumerate([m2.grad] + [t[k].grad for k in names]):
        to_host(4 + j, g)
    ev["grad_d2h_done"].record(main); ht["grad_copies_issued"] = time.perf_counter()
    main.synchronize(); side.synchronize()
    ht["synced"] = time.perf_counter()
    return ({k: ev["start"].elapsed_time(ev[k]) * 1e3 for k in EV[1:]}, {k: (ht[k] - t0) * 1e6 for k in HT})


for _ in range(10):
    step()
acc_e, acc_h, n = {}, {}, 100
for _ in range(n):
    flush.zero_(); torch.cuda.synchronize()
    e, h = step()
    for k, v in e.items(): acc_e[k] = acc_e.get(k, 0.0) + v / n
    for k, v in h.items(): acc_h[k] = acc_h.get(k, 0.0) + v / n
print("GPU  (us after start):", {k: round(v, 1) for k, v in acc_e.items()})
print("host (us after start):", {k: round(v, 1) for k, v in acc_h.items()})
