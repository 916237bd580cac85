"""GPU: NVTX ranges (moss_nvtx_enable / MOSS_NVTX=1) change nothing but the annotations: a step with
the ranges on gives the bits of a step with them off, inside and outside a profiler.  What the ranges look like to a tool is
checked by tools/ncu_nvtx.sh (`ncu --nvtx --nvtx-include "moss_rasterize_backward/"` keeps exactly the backward's launches;
its output is under profiles/)."""
import pytest
import torch

import moss_b200 as M
from moss_b200 import _abi, synth
from tests import scenes as S

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def test_ranges_do_not_change_results():
    lib = _abi.load()
    d = S.make_scene("small")
    cam = d["cam"]
    rs = M.GaussianRasterizationSettings(**synth.settings_kwargs(cam, d["bg"], d["sh_degree"], d["scale_modifier"], device=DEV))
    keys = ("means3D", "shs", "colors_precomp", "opacities", "scales", "rotations", "cov3D_precomp")
    t = [d[k].to(DEV) if d[k] is not None else torch.empty(0) for k in keys]
    grads = tuple(g.to(DEV) for g in S.upstream_grads(cam.image_width, cam.image_height))

    def step():
        c, r, dp, a, st = M.forward_raw(*t, rs)
        g = M.backward_raw(st, r, *grads, flags=_abi.FLAG_DETERMINISTIC)
        torch.cuda.synchronize()
        return [c, r, dp, a, *g]

    off = step()
    assert lib.moss_nvtx_enable(1) == 0
    try:
        launches0 = lib.moss_kernel_launches()
        on = step()
        assert lib.moss_kernel_launches() - launches0 >= 7        # the same launches, each inside its range
    finally:
        assert lib.moss_nvtx_enable(0) == 0
    for a, b in zip(off, on):
        assert torch.equal(a, b)
