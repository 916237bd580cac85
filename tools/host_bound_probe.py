"""Developer tool: is the ZJU step GPU-bound or host-bound?  Wall time per fwd+bwd for the real scene vs a tiny
scene (host floor), synchronous vs MOSS_FLAG_ASYNC."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import moss_b200 as M
from moss_b200 import synth, _abi
from tests import scenes as S

dev = torch.device("cuda:0")
for name in ("tiny", "zju"):
    d = S.make_scene(name)
    cam = d["cam"]
    rs = M.GaussianRasterizationSettings(**synth.settings_kwargs(cam, d["bg"], d["sh_degree"], d["scale_modifier"], device=dev))
    t = {k: (d[k].to(dev) if d[k] is not None else torch.empty(0)) for k in ("means3D", "shs", "colors_precomp", "opacities", "scales", "rotations", "cov3D_precomp")}
    gc, gd, ga = (x.to(dev) for x in S.upstream_grads(cam.image_width, cam.image_height, depth=False))
    for flags, label in ((0, "sync"), (_abi.FLAG_ASYNC, "async")):
        def step():
            c, r, dp, a, st = M.forward_raw(t["means3D"], t["shs"], t["colors_precomp"], t["opacities"], t["scales"], t["rotations"], t["cov3D_precomp"], rs, flags=flags)
            M.backward_raw(st, r, gc, gd, ga)
        for _ in range(20): step()
        torch.cuda.synchronize()
        n = 300
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter(); e0.record()
        for _ in range(n): step()
        t_issue = time.perf_counter() - t0
        e1.record(); torch.cuda.synchronize()
        t_all = time.perf_counter() - t0
        print(f"{name:5s} {label:5s}: host issue {t_issue/n*1e6:7.1f} us/step, wall {t_all/n*1e6:7.1f} us/step, gpu {e0.elapsed_time(e1)/n*1e3:7.1f} us/step")
