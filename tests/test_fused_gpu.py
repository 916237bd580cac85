"""GPU: the fused skin + rasterize entry points (moss_skin_rasterize_forward / _backward; north_star subsystem 1) by
SURVEY.md Appendix A.6's two-step procedure:
  1. the skinned centres / transforms the fused kernel computes in registers (optional outputs) against the front-end
     kernel (itself pinned on goldens of the REAL reference python functions, tests/test_skin_gpu.py): bit for bit;
  2. the fused call's images, radii and binning against the UNMODIFIED reference rasterizer (oracle/_ref) fed with those
     centres and the front end's covariance: bit for bit -- the comparison the rasterizer rows a7-a12 are held to;
and its gradients (parameters, frame tables, SH, opacity, screen-space sink) against the two-call form, whose two
kernels are each pinned on the reference (tests/test_parity_gpu.py, tests/test_skin_gpu.py): <= 1e-5 norm-relative
(identical arithmetic up to the order of the atomics)."""
import pytest
import torch

import moss_b200 as M
from moss_b200 import fused as F, skinning as K, synth
from tests import scenes as S
from tests.lbs_cases import oracle_smpl

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _setup(P, W, H, seed, with_delta, cam_index=2):
    from oracle import lbs_oracle as L
    body = synth.body()
    g = synth.gaussians(P, body, seed=synth.SEED + seed)
    fr = synth.frame_params(seed)
    fr.params["Th"] = torch.zeros(1, 3); fr.params["R"] = torch.eye(3)       # keep the body in front of the ring camera
    ft = K.FrameTables(*[t.to(DEV).contiguous() for t in L.frame_tables(oracle_smpl(body), fr.params, fr.big_params)])
    cam = synth.ring_cameras(23, W, H)[cam_index]
    bg = torch.tensor([0.1, 0.2, 0.3])
    rs = M.GaussianRasterizationSettings(**synth.settings_kwargs(cam, bg, 3, 1.0, device=DEV))
    gen = torch.Generator().manual_seed(seed)
    delta = (0.3 * torch.randn(P, 24, generator=gen)).to(DEV) if with_delta else None
    raw_rot = g.rotations * (0.7 + 0.7 * torch.rand(P, 1, generator=gen))          # stored quaternions are not unit
    return body, g, ft, cam, bg, rs, delta, raw_rot


def _rel(a, b):
    a, b = a.detach().double(), b.detach().double()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


CASES = [(20000, 512, 512, 9, False), (5001, 200, 120, 4, True), (129, 64, 48, 5, True), (100000, 1024, 1024, 3, True)]


@pytest.mark.parametrize("P,W,H,seed,with_delta", CASES)
def test_forward_two_step_parity(P, W, H, seed, with_delta):
    body, g, ft, cam, bg, rs, delta, raw_rot = _setup(P, W, H, seed, with_delta)
    xyz, sc, vid, wts = g.xyz.to(DEV), g.scales.to(DEV), g.vert_ids.to(DEV), body.weights.to(DEV)
    rot, shs, opac = raw_rot.to(DEV), g.shs.to(DEV), g.opacities.to(DEV)
    # the two-call form
    m, cov, tr, _, _ = K.skin_and_covariance(xyz, vid, wts, ft, sc, rot, 1.0, delta)
    c2, r2, d2, a2 = M.GaussianRasterizer(rs)(means3D=m, means2D=torch.zeros_like(m), shs=shs, colors_precomp=None,
                                              opacities=opac, scales=None, rotations=None, cov3D_precomp=cov)
    # the fused call, with the optional per-Gaussian outputs through the raw hook
    k = K._fill_skin(xyz, vid.to(torch.int32), wts, delta, ft.A_big, ft.A_tgt, ft.poseoff_big, ft.shapeoff, ft.poseoff_tgt,
                     ft.R, ft.Th, sc, rot, 1.0)
    m_f, tr_f = torch.empty_like(m), torch.empty_like(tr)
    from moss_b200 import rasterizer as R
    c1, r1, d1, a1, st = R._forward(xyz, shs, R._EMPTY, opac, R._EMPTY, R._EMPTY, R._EMPTY, [rs], 0, False,
                                    skin=(k, (), m_f, tr_f))
    # step 1: what the fused kernel held in registers == what the front-end kernel wrote
    assert torch.equal(m_f, m), float((m_f - m).abs().max())
    assert torch.equal(tr_f, tr)
    # the two-call form and the fused call agree bit for bit
    assert torch.equal(r1, r2) and int((r1 > 0).sum()) > P // 3
    for x, y in ((c1, c2), (d1, d2), (a1, a2)):
        assert torch.equal(x, y), float((x - y).abs().max())
    # the public wrapper returns the same
    c3, r3, d3, a3, tr3 = F.skin_and_rasterize(rs, xyz, vid, wts, ft, sc, rot, opac, shs=shs, lbs_delta=delta)
    assert torch.equal(c3, c1) and torch.equal(r3, r1) and torch.equal(d3, d1) and torch.equal(a3, a1) and torch.equal(tr3, tr)
    # step 2: the unmodified reference rasterizer on the fused centres + the front end's covariance
    from oracle import ref_shim
    if not ref_shim.available():
        pytest.skip("oracle/_ref/libref_raster.so not built (python oracle/build_ref.py shim)")
    d = dict(means3D=m_f, shs=shs, colors_precomp=None, opacities=opac, scales=None, rotations=None, cov3D_precomp=cov,
             cam=cam, bg=bg, sh_degree=3, scale_modifier=1.0)
    ref = ref_shim.RefCall(d)
    rc, rr, rd, ra = ref.forward()
    assert torch.equal(r1, rr)
    assert st.num_rendered == ref.R
    for nm, x, y in (("color", c1, rc), ("depth", d1, rd), ("alpha", a1, ra)):
        assert torch.equal(x, y), f"{nm}: {float((x - y).abs().max())}"
    assert torch.equal(M.export_state(st, "ranges"), ref.state("ranges"))
    assert torch.equal(M.export_state(st, "n_contrib"), ref.state("n_contrib"))


@pytest.mark.parametrize("P,W,H,seed,with_delta", CASES)
def test_backward_matches_two_call_form(P, W, H, seed, with_delta):
    body, g, ft, cam, bg, rs, delta, raw_rot = _setup(P, W, H, seed, with_delta)
    vid, wts = g.vert_ids.to(DEV), body.weights.to(DEV)
    gc, gd, ga = (t.to(DEV) for t in S.upstream_grads(W, H))

    def leaves():
        leaf = lambda t: t.to(DEV).clone().requires_grad_(True)  # noqa: E731
        d = dict(xyz=leaf(g.xyz), sc=leaf(g.scales), rot=leaf(raw_rot), shs=leaf(g.shs), opac=leaf(g.opacities),
                 A=leaf(ft.A_tgt), po=leaf(ft.poseoff_tgt), m2d=torch.zeros(P, 3, device=DEV, requires_grad=True))
        d["delta"] = leaf(delta) if delta is not None else None
        return d
    a = leaves()
    fta = ft._replace(A_tgt=a["A"], poseoff_tgt=a["po"])
    m, cov, _, _, _ = K.skin_and_covariance(a["xyz"], vid, wts, fta, a["sc"], a["rot"], 1.0, a["delta"])
    c, r, dp, al = M.GaussianRasterizer(rs)(means3D=m, means2D=a["m2d"], shs=a["shs"], colors_precomp=None, opacities=a["opac"],
                                            scales=None, rotations=None, cov3D_precomp=cov)
    torch.autograd.backward([c, dp, al], [gc, gd, ga])
    b = leaves()
    ftb = ft._replace(A_tgt=b["A"], poseoff_tgt=b["po"])
    c1, r1, d1, a1, _ = F.skin_and_rasterize(rs, b["xyz"], vid, wts, ftb, b["sc"], b["rot"], b["opac"], means2D=b["m2d"],
                                             shs=b["shs"], lbs_delta=b["delta"])
    torch.autograd.backward([c1, d1, a1], [gc, gd, ga])
    for key in ("xyz", "sc", "rot", "shs", "opac", "A", "po", "m2d") + (("delta",) if delta is not None else ()):
        assert b[key].grad is not None and b[key].grad.shape == a[key].grad.shape, key
        assert _rel(b[key].grad, a[key].grad) < 1e-5, (key, _rel(b[key].grad, a[key].grad))
    assert float(a["xyz"].grad.abs().sum()) > 0 and float(a["A"].grad.abs().sum()) > 0


def test_colors_precomp_absent_upstreams_and_argument_errors():
    P, W, H = 2000, 96, 80
    body, g, ft, cam, bg, rs, delta, raw_rot = _setup(P, W, H, 6, False)
    vid, wts = g.vert_ids.to(DEV), body.weights.to(DEV)
    leaf = lambda t: t.to(DEV).clone().requires_grad_(True)  # noqa: E731
    col = leaf(torch.rand(P, 3, generator=torch.Generator().manual_seed(1)))
    xyz, sc, rot, opac = leaf(g.xyz), leaf(g.scales), leaf(raw_rot), leaf(g.opacities)
    c, r, d, a, tr = F.skin_and_rasterize(rs, xyz, vid, wts, ft, sc, rot, opac, colors_precomp=col, return_transforms=False)
    assert tr is None
    c.sum().backward()                               # depth / alpha receive no gradient: NULL upstreams
    m, cov, _, _, _ = K.skin_and_covariance(g.xyz.to(DEV), vid, wts, ft, g.scales.to(DEV), raw_rot.to(DEV), 1.0)
    col2 = col.detach().clone().requires_grad_(True)
    c2, _, _, _ = M.GaussianRasterizer(rs)(means3D=m, means2D=torch.zeros_like(m), shs=None, colors_precomp=col2,
                                           opacities=g.opacities.to(DEV), scales=None, rotations=None, cov3D_precomp=cov)
    c2.sum().backward()
    assert torch.equal(c, c2) and _rel(col.grad, col2.grad) < 1e-5 and float(xyz.grad.abs().sum()) > 0
    with pytest.raises(Exception, match="excatly one"):
        F.skin_and_rasterize(rs, xyz, vid, wts, ft, sc, rot, opac)
    with pytest.raises(RuntimeError, match="scaling and rotation"):
        F.skin_and_rasterize(rs, xyz, vid, wts, ft, None, None, opac, colors_precomp=col)
    with pytest.raises(RuntimeError, match="no CPU path"):
        F.skin_and_rasterize(rs, g.xyz, vid, wts, ft, sc, rot, opac, colors_precomp=col)
