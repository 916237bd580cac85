"""Developer tool: fwd+bwd captured in a CUDA graph (MOSS_FLAG_ASYNC) vs stream launches; warm and L2-flushed."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import moss_b200 as M
from moss_b200 import synth, _abi
from tests import scenes as S

dev = torch.device("cuda:0")
flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=dev)
for name in sys.argv[1:] or ("tiny", "zju"):
    d = S.make_scene(name)
    cam = d["cam"]
    rs = M.GaussianRasterizationSettings(**synth.settings_kwargs(cam, d["bg"], d["sh_degree"], d["scale_modifier"], device=dev))
    t = {k: (d[k].to(dev) if d[k] is not None else torch.empty(0)) for k in ("means3D", "shs", "colors_precomp", "opacities", "scales", "rotations", "cov3D_precomp")}
    gc, gd, ga = (x.to(dev) for x in S.upstream_grads(cam.image_width, cam.image_height, depth=False))
    def step(flags):
        c, r, dp, a, st = M.forward_raw(t["means3D"], t["shs"], t["colors_precomp"], t["opacities"], t["scales"], t["rotations"], t["cov3D_precomp"], rs, flags=flags)
        M.backward_raw(st, r, gc, gd, ga)
        return st
    for _ in range(5): step(0)            # learn the capacity hint
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(3): step(_abi.FLAG_ASYNC)
    torch.cuda.current_stream().wait_stream(s)
    with torch.cuda.graph(g):
        st = step(_abi.FLAG_ASYNC)
    def timeit(fn, do_flush, n=200):
        for _ in range(10): fn()
        torch.cuda.synchronize()
        tot = 0.0
        evs = []
        for _ in range(n):
            if do_flush: flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record(); evs.append((e0, e1))
        torch.cuda.synchronize()
        return sum(a.elapsed_time(b) for a, b in evs) / n * 1e3
    for do_flush in (False, True):
        print(f"{name:5s} flush={do_flush}: stream sync {timeit(lambda: step(0), do_flush):7.1f} us  stream async {timeit(lambda: step(_abi.FLAG_ASYNC), do_flush):7.1f} us  graph {timeit(g.replay, do_flush):7.1f} us")
