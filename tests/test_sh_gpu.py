"""GPU: the kernel's SH -> colour path (csrc/preprocess.cuh::sh_to_rgb, inside K1) against the reference's python
``eval_sh`` (utils/sh_utils.py:57-112) through the committed golden file tests/golden/eval_sh.npz:
  * the per-Gaussian colours K1 stores (state export) equal ``clamp_min(eval_sh + 0.5, 0)`` to 2e-6 (a few ulps);
  * rendering with ``shs`` equals rendering with ``colors_precomp = that golden colour`` -- the reference's
    ``convert_SHs_python`` switch (gaussian_renderer/__init__.py:100-105) -- to 1e-5 max-abs on the image, with identical
    radii / instance lists (colours do not enter the alpha thresholds);
  * markVisible equals the unmodified reference's (bool-exact), including points on the z = 0.2 plane (row a17)."""
import os

import numpy as np
import pytest
import torch

import moss_b200 as M
from moss_b200 import _abi, synth
from tests import scenes as S

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "eval_sh.npz")
DEV = "cuda"


def _run(d, colors=None):
    to = lambda t: torch.empty(0) if t is None else t.to(DEV)  # noqa: E731
    rs = M.GaussianRasterizationSettings(**synth.settings_kwargs(d["cam"], d["bg"], d["sh_degree"], d["scale_modifier"], device=DEV))
    shs = None if colors is not None else d["shs"]
    return M.forward_raw(to(d["means3D"]), to(shs), to(colors), to(d["opacities"]), to(d["scales"]), to(d["rotations"]),
                         to(d["cov3D_precomp"]), rs, flags=_abi.FLAG_EXPORT_KEYS)


@pytest.mark.parametrize("name", ["small", "small_d0", "small_d1", "small_d2", "ragged"])
def test_kernel_sh_equals_reference_eval_sh(name):
    z = np.load(GOLDEN)
    d = S.make_scene(name)
    gold = torch.from_numpy(z[name])
    color, radii, depth, alpha, st = _run(d)
    vis = (radii > 0).cpu()
    rgb = M.export_state(st, "rgb").cpu()
    assert int(vis.sum()) > 100
    assert float((rgb[vis] - gold[vis]).abs().max()) <= 2e-6
    # the clamp flags agree wherever the reference's value is not within rounding of zero
    cl = M.export_state(st, "clamped").cpu().bool()
    raw = torch.from_numpy(z[name + "_raw"])
    sure = raw.abs() > 2e-6
    assert torch.equal(cl[vis][sure[vis]], (raw < 0)[vis][sure[vis]])
    c2, r2, d2, a2, st2 = _run(d, colors=gold)
    assert torch.equal(radii, r2)
    assert torch.equal(M.export_state(st, "point_list"), M.export_state(st2, "point_list"))
    assert torch.equal(alpha, a2) and torch.equal(depth, d2)
    assert float((color - c2).abs().max()) <= 1e-5


def test_mark_visible_equals_reference():
    from oracle import ref_shim
    if not ref_shim.available():
        pytest.skip("oracle/_ref/libref_raster.so not built")
    lib = ref_shim.load()
    for name in ("ragged", "small", "zju"):
        d = S.make_scene(name)
        cam = d["cam"]
        pts = d["means3D"].clone()
        # put a few hundred points exactly on, just in front of and just behind the z = 0.2 culling plane
        W2C = cam.world_view_transform.t()            # column-vector form: p_view = W2C @ [p,1]
        Rm, tv = W2C[:3, :3], W2C[:3, 3]
        n = min(300, pts.shape[0])
        for k, z0 in enumerate((0.2, 0.2000001, 0.1999999, 0.0, -1.0)):
            pv = torch.tensor([0.05 * k, -0.03 * k, z0]).repeat(n // 5, 1)
            pv[:, :2] += torch.linspace(-0.5, 0.5, n // 5)[:, None]
            pts[k * (n // 5):(k + 1) * (n // 5)] = (pv - tv) @ Rm     # R^T (pv - t), written for row vectors
        pts = pts.to(DEV).contiguous()
        rs = M.GaussianRasterizationSettings(**synth.settings_kwargs(cam, d["bg"], d["sh_degree"], 1.0, device=DEV))
        ours = M.GaussianRasterizer(rs).markVisible(pts)
        ref = torch.zeros(pts.shape[0], dtype=torch.bool, device=DEV)
        view = cam.world_view_transform.to(DEV).contiguous()
        proj = cam.full_proj_transform.to(DEV).contiguous()
        torch.cuda.synchronize()
        lib.ref_mark_visible(pts.shape[0], pts.data_ptr(), view.data_ptr(), proj.data_ptr(), ref.data_ptr())
        torch.cuda.synchronize()
        assert ours.dtype == torch.bool and torch.equal(ours, ref), f"{name}: {(ours != ref).sum().item()} differ"
        assert 0 < int(ref.sum()) < pts.shape[0]
