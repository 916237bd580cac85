"""GPU: the fused skinning + covariance kernel (through the C ABI) against
  * tests/golden/lbs_*.npz -- outputs AND autograd gradients of the real reference python functions,
  * the CPU oracle (oracle/lbs_oracle.py) on seeded inputs at ZJU size, including the table gradients,
  * and, end to end, skin -> rasterizer against MOSS's cov3D_precomp path fed with the oracle's skinning.
Tolerances (fp32, different summation order than torch's batched matmuls / LU inverse): forward 1e-5 relative to the
tensor's max-abs, gradients 1e-4 norm-relative (north_star allows 1e-3)."""
import pytest
import torch

from moss_b200 import skinning as K, synth
from tests.lbs_cases import LBS_CASES, load_case, oracle_smpl

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
FWD_TOL = 1e-5
GRAD_TOL = 1e-4


def _close(a, b, tol=FWD_TOL):
    a, b = a.detach().cpu().double(), b.detach().cpu().double()
    scale = max(float(b.abs().max()), 1.0)
    err = float((a - b).abs().max()) / scale
    assert err <= tol, err


def _rel(a, b):
    a, b = a.detach().cpu().double(), b.detach().cpu().double()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def _to(ft, dev):
    return K.FrameTables(*[t.to(dev).contiguous() for t in ft])


@pytest.mark.parametrize("name", LBS_CASES)
def test_forward_matches_reference_golden(name):
    d, ft, body = load_case(name)
    delta = d["lbs_delta"].to(DEV) if "lbs_delta" in d else None
    m, cov, tr, bw, tl = K.skin_and_covariance(d["xyz"].to(DEV), d["vert_ids"].to(DEV), body.weights.to(DEV), _to(ft, DEV),
                                               d["scaling"].to(DEV), d["rotation"].to(DEV), float(d["mod"]), delta,
                                               return_transl=True)
    _close(m, d["world"]); _close(tr, d["transforms"]); _close(bw, d["bweights"]); _close(tl, d["translation"])
    assert _rel(cov, d["cov3D"]) < 1e-5
    _close(cov, d["cov3D"], 2e-5 * float(d["cov3D"].abs().max()) / max(float(d["cov3D"].abs().max()), 1.0) + 1e-9)


@pytest.mark.parametrize("name", LBS_CASES)
def test_backward_matches_reference_golden(name):
    d, ft, body = load_case(name)
    leaf = lambda t: t.to(DEV).requires_grad_(True)  # noqa: E731
    xyz, sc, rot = leaf(d["xyz"]), leaf(d["scaling"]), leaf(d["rotation"])
    delta = leaf(d["lbs_delta"]) if "lbs_delta" in d else None
    m, cov, _, _, _ = K.skin_and_covariance(xyz, d["vert_ids"].to(DEV), body.weights.to(DEV), _to(ft, DEV), sc, rot,
                                            float(d["mod"]), delta)
    torch.autograd.backward([m, cov], [d["g_world"].to(DEV), d["g_cov"].to(DEV)])
    assert _rel(xyz.grad, d["d_xyz"]) < GRAD_TOL
    assert _rel(sc.grad, d["d_scaling"]) < GRAD_TOL
    assert _rel(rot.grad, d["d_rotation"]) < GRAD_TOL
    if delta is not None:
        assert _rel(delta.grad, d["d_lbs_delta"]) < GRAD_TOL


@pytest.mark.parametrize("P,with_delta", [(20000, False), (20000, True), (100000, True), (5, True), (129, False)])
def test_against_oracle_with_table_gradients(P, with_delta):
    """ZJU / MonoCap sizes, ragged tails, gradients w.r.t. the target joint transforms and pose offsets (the path
    correct_Rs trains through, scene/gaussian_model.py:880-895)."""
    from oracle import lbs_oracle as L
    body = synth.body()
    g = synth.gaussians(P, body, seed=synth.SEED + 5)
    fr = synth.frame_params(7)
    ft = L.frame_tables(oracle_smpl(body), fr.params, fr.big_params)
    gen = torch.Generator().manual_seed(P)
    delta = 0.3 * torch.randn(P, 24, generator=gen) if with_delta else None
    gw, gc = torch.randn(P, 3, generator=gen), 50.0 * torch.randn(P, 6, generator=gen)
    # oracle (CPU autograd)
    A_t = ft.A_tgt.clone().requires_grad_(True)
    po_t = ft.poseoff_tgt.clone().requires_grad_(True)
    xyz_c, sc_c, rot_c = (t.clone().requires_grad_(True) for t in (g.xyz, g.scales, g.rotations))
    del_c = delta.clone().requires_grad_(True) if with_delta else None
    sk, cov_c = L.skin_and_covariance(xyz_c, g.vert_ids, body.weights, ft._replace(A_tgt=A_t, poseoff_tgt=po_t), sc_c,
                                      rot_c, 1.0, del_c)
    torch.autograd.backward([sk.world_pts, cov_c], [gw, gc])
    # kernel
    leaf = lambda t: t.to(DEV).requires_grad_(True)  # noqa: E731
    ftd = _to(ft, DEV)
    A_d, po_d = leaf(ft.A_tgt), leaf(ft.poseoff_tgt)
    xyz, sc, rot = leaf(g.xyz), leaf(g.scales), leaf(g.rotations)
    dl = leaf(delta) if with_delta else None
    m, cov, tr, bw, _ = K.skin_and_covariance(xyz, g.vert_ids.to(DEV), body.weights.to(DEV),
                                              ftd._replace(A_tgt=A_d, poseoff_tgt=po_d), sc, rot, 1.0, dl)
    torch.autograd.backward([m, cov], [gw.to(DEV), gc.to(DEV)])
    _close(m, sk.world_pts); _close(tr, sk.transforms); _close(bw, sk.bweights)
    assert _rel(cov, cov_c) < 1e-5
    assert _rel(xyz.grad, xyz_c.grad) < GRAD_TOL
    assert _rel(sc.grad, sc_c.grad) < GRAD_TOL
    assert _rel(rot.grad, rot_c.grad) < GRAD_TOL
    assert _rel(A_d.grad[:, :3], A_t.grad[:, :3]) < GRAD_TOL
    assert float(A_d.grad[:, 3].abs().max()) == 0.0
    assert _rel(po_d.grad, po_t.grad) < GRAD_TOL
    if with_delta:
        assert _rel(dl.grad, del_c.grad) < GRAD_TOL


def test_skin_then_rasterize_end_to_end():
    """render() as MOSS runs it (gaussian_renderer/__init__.py:57-120): skinning -> cov3D_precomp -> rasterizer, with
    gradients flowing back to xyz / scaling / rotation.  Checked against the oracle chain (LBS oracle + raster oracle)."""
    import moss_b200 as M
    from oracle import lbs_oracle as L, raster_oracle as O
    from tests import scenes as S
    body = synth.body()
    P, W, H = 3000, 128, 128
    g = synth.gaussians(P, body, seed=synth.SEED + 9)
    fr = synth.frame_params(11)
    fr.params["Th"] = torch.zeros(1, 3); fr.params["R"] = torch.eye(3)       # keep the body in front of the ring camera
    ft = L.frame_tables(oracle_smpl(body), fr.params, fr.big_params)
    cam = synth.ring_cameras(23, W, H)[2]
    bg = torch.tensor([0.1, 0.2, 0.3])
    # oracle chain
    xyz_c, sc_c, rot_c = (t.clone().requires_grad_(True) for t in (g.xyz, g.scales, g.rotations))
    sk, cov_c = L.skin_and_covariance(xyz_c, g.vert_ids, body.weights, ft, sc_c, rot_c, 1.0)
    s = O.Settings(**synth.settings_kwargs(cam, bg, 3, 1.0))
    fw = O.rasterize_forward(sk.world_pts.detach(), g.shs, None, g.opacities, None, None, cov_c.detach(), s)
    gc, gd, ga = S.upstream_grads(W, H)
    og = O.rasterize_backward(fw, sk.world_pts.detach(), g.shs, None, None, None, cov_c.detach(), s, gc, gd, ga)
    torch.autograd.backward([sk.world_pts, cov_c], [og.means3D.float(), og.cov3D_precomp.float()])
    # product chain
    leaf = lambda t: t.to(DEV).requires_grad_(True)  # noqa: E731
    xyz, sc, rot = leaf(g.xyz), leaf(g.scales), leaf(g.rotations)
    m, cov, _, _, _ = K.skin_and_covariance(xyz, g.vert_ids.to(DEV), body.weights.to(DEV), _to(ft, DEV), sc, rot, 1.0)
    rs = M.GaussianRasterizationSettings(**synth.settings_kwargs(cam, bg, 3, 1.0, device=DEV))
    means2D = torch.zeros_like(m, requires_grad=True)
    color, radii, depth, alpha = M.GaussianRasterizer(rs)(means3D=m, means2D=means2D, shs=g.shs.to(DEV), colors_precomp=None,
                                                          opacities=g.opacities.to(DEV), scales=None, rotations=None,
                                                          cov3D_precomp=cov)
    torch.autograd.backward([color, depth, alpha], [gc.to(DEV), gd.to(DEV), ga.to(DEV)])
    assert int((radii > 0).sum()) > P // 2
    assert float((color.cpu() - fw.color).abs().max()) < 2e-3      # skinning differs by ~1e-7 -> a few splats flip a pixel test
    assert float((color.cpu() - fw.color).abs().mean()) < 1e-5
    for a, b in ((xyz.grad, xyz_c.grad), (sc.grad, sc_c.grad), (rot.grad, rot_c.grad)):
        assert _rel(a, b) < 2e-2      # pixel-test flips dominate; tight checks are in the two tests above


# ---- nearest SMPL vertex (replaces the knn_cuda call at scene/gaussian_model.py:827) ---------------------------------
def _check_nearest(q, v, ids, dist=None):
    """ids must attain the minimum distance (fp64), and where the fp64 argmin differs the two candidates are tied to
    fp32 resolution; on exact ties the lowest index wins."""
    from oracle import lbs_oracle as L
    ref = L.nearest_vertex(v.cpu(), q.cpu())
    ids = ids.cpu().long()
    assert ids.min() >= 0 and ids.max() < v.shape[0]
    dd = torch.cdist(q.cpu().double(), v.cpu().double())
    d_ours = dd.gather(1, ids[:, None])[:, 0]
    d_ref = dd.gather(1, ref[:, None])[:, 0]
    differ = ids != ref
    assert float(differ.float().mean()) < 1e-3
    assert torch.all(d_ours[differ] <= d_ref[differ] * (1 + 1e-6) + 1e-12)
    if dist is not None:
        assert torch.allclose(dist.cpu().double(), d_ours, rtol=1e-6, atol=1e-9)


@pytest.mark.parametrize("name", LBS_CASES)
def test_nearest_vertex_matches_reference_golden(name):
    """The golden files hold the vert_ids the REAL coarse_deform_c2source used (exact k=1 search)."""
    d, ft, body = load_case(name)
    ids = K.nearest_vertex(d["xyz"].to(DEV)[None], body.verts.to(DEV)[None])
    assert ids.dtype == torch.int32
    assert torch.equal(ids.cpu(), d["vert_ids"].to(torch.int32).cpu())


@pytest.mark.parametrize("P,V", [(1, 1), (129, 1023), (1000, 1025), (20000, 6890), (5, 2500)])
def test_nearest_vertex_matches_oracle(P, V):
    g = torch.Generator().manual_seed(P * 7919 + V)
    v = torch.randn(V, 3, generator=g)
    q = v[torch.randint(0, V, (P,), generator=g)] + 0.05 * torch.randn(P, 3, generator=g)
    ids, dist = K.nearest_vertex(q.to(DEV), v.to(DEV), return_distance=True)
    _check_nearest(q, v, ids, dist)


def test_nearest_vertex_ties_take_the_lowest_index():
    v = torch.tensor([[0., 0., 0.], [1., 0., 0.], [1., 0., 0.], [1., 0., 0.], [1., 0., 0.], [1., 0., 0.], [-1., 0., 0.]])
    v = torch.cat([v, v + 10.0])                           # 14 vertices: duplicates across the 4-way unroll
    q = torch.tensor([[1., 0., 0.], [0.9, 0., 0.], [0., 0., 0.], [11., 10., 10.], [-1.2, 0., 0.]])
    ids = K.nearest_vertex(q.to(DEV), v.to(DEV)).cpu().tolist()
    assert ids == [1, 1, 0, 8, 6]
    # equidistant from vertex 0 and vertex 1 (exactly representable): lowest index
    assert K.nearest_vertex(torch.tensor([[0.5, 0., 0.]], device=DEV), v.to(DEV)).item() == 0


# ---- the grid search gives the exhaustive search's answers ---------------------------------------------------------------
def _both(q, v):
    qd, vd = q.to(DEV), v.to(DEV)
    a, da = K.nearest_vertex(qd, vd, return_distance=True, grid=None)
    g = K.VertexGrid(vd)
    b, db = K.nearest_vertex(qd, vd, return_distance=True, grid=g)
    assert torch.equal(a, b), int((a != b).sum())
    assert torch.equal(da, db)
    return a


@pytest.mark.parametrize("P,V,spread", [(20000, 6890, 0.03), (100000, 6890, 0.03), (5000, 6890, 2.0), (3000, 64, 0.5), (1000, 1, 1.0),
                                        (4097, 1025, 0.0)])
def test_grid_search_equals_exhaustive_search(P, V, spread):
    """Near-surface queries (what training sees), far-away queries (rings out to the whole grid), a grid of one vertex,
    and queries ON the vertices (zero distance, the lowest index among duplicates)."""
    body = synth.body()
    g = torch.Generator().manual_seed(P + 31 * V)
    v = body.verts[torch.randperm(body.verts.shape[0], generator=g)[:V]].clone() if V <= body.verts.shape[0] else torch.randn(V, 3, generator=g)
    if V >= 1025:
        v[V // 2:V // 2 + 50] = v[:50]                          # duplicated vertices: ties by index
    q = v[torch.randint(0, V, (P,), generator=g)] + spread * torch.randn(P, 3, generator=g)
    ids = _both(q, v)
    if P * V <= 150_000_000:            # the fp64 yardstick builds a P x V matrix on the host
        _check_nearest(q, v, ids)


def test_grid_search_degenerate_sets_and_auto_cache():
    # all vertices in one plane / on one point; NaN and infinite queries
    v = torch.zeros(300, 3); v[:, 0] = torch.linspace(-1, 1, 300)
    q = torch.randn(500, 3)
    q[0] = float("nan"); q[1, 2] = float("inf"); q[2] = torch.tensor([1e20, -1e20, 0.0])
    _both(q, v)
    _both(torch.randn(100, 3), torch.ones(70, 3))
    # the default path builds the grid once per vertex tensor and rebuilds it after an in-place edit
    body = synth.body()
    vd = body.verts.to(DEV)
    qd = (body.verts[:4000] + 0.02 * torch.randn(4000, 3)).to(DEV)
    a = K.nearest_vertex(qd, vd)
    g1 = K._grid_for(vd)
    assert K._grid_for(vd) is g1
    assert torch.equal(a, K.nearest_vertex(qd, vd, grid=None))
    vd.add_(0.5)
    assert K._grid_for(vd) is not g1
    assert torch.equal(K.nearest_vertex(qd, vd), K.nearest_vertex(qd, vd, grid=None))
    with pytest.raises(RuntimeError, match="another vertex set"):
        K.nearest_vertex(qd, vd[:100], grid=g1)
