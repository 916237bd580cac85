"""GPU: batched multi-view rasterization (moss_rasterize_forward_batch / backward_batch; SURVEY.md section 7 step 7,
BASELINE.json config 4) -- V cameras of the same Gaussians in ONE launch sequence.

Per view the batch must reproduce what V separate calls of the UNMODIFIED reference produce on the same tensors
(train_ZJU.py:93-100 renders one camera per iteration): radii, tile counts, sorted key / ID lists, ranges, n_contrib,
final_T and the image bit for bit; the parameter gradients are the sum over the views of the reference's gradients
(<= 1e-3 relative, north_star), the screen-space gradient stays per view."""
import pytest
import torch

import moss_b200 as M
from moss_b200 import _abi, synth
from tests import scenes as S
from tests.test_parity_gpu import rel_err, _bits, GRAD_TOL

pytestmark = pytest.mark.gpu
DEV = "cuda"
NAMES = ("means2D", "colors", "opacity", "means3D", "cov3D", "sh", "scales", "rotations")


@pytest.fixture(scope="module")
def ref_mod():
    from oracle import ref_shim
    if not ref_shim.available():
        pytest.skip("oracle/_ref/libref_raster.so not built (python oracle/build_ref.py shim)")
    return ref_shim


def _cams(d, n_views, step=1):
    cam = d["cam"]
    ring = synth.ring_cameras(23, cam.image_width, cam.image_height)
    return [ring[(k * step) % 23] for k in range(n_views)]


def _settings(d, cam):
    return M.GaussianRasterizationSettings(**synth.settings_kwargs(cam, d["bg"], d["sh_degree"], d["scale_modifier"], device=DEV))


def _dev(d):
    to = lambda t: torch.empty(0) if t is None else t.to(DEV)  # noqa: E731
    return [to(d[k]) for k in ("means3D", "shs", "colors_precomp", "opacities", "scales", "rotations", "cov3D_precomp")]


def _upstream(V, W, H, depth=True):
    gs = [S.upstream_grads(W, H, seed=7 + v, depth=depth) for v in range(V)]
    return tuple(torch.stack([g[i] for g in gs]).to(DEV) for i in range(3))


def run_batch(d, cams, flags=0, grads=None, bflags=0):
    rs = [_settings(d, c) for c in cams]
    color, radii, depth, alpha, st = M.forward_batch_raw(*_dev(d), rs, flags=flags)
    out = dict(color=color, radii=radii, depth=depth, alpha=alpha, st=st)
    if grads is not None:
        out["grads"] = dict(zip(NAMES, M.backward_raw(st, radii, *grads, flags=bflags)))
    torch.cuda.synchronize()
    return out


@pytest.mark.parametrize("name,V,step", [("zju", 23, 1), ("small", 5, 4), ("ragged", 3, 7), ("small_cov", 2, 11),
                                          ("small_color", 4, 5), ("big_splats", 6, 3), ("tiny", 1, 1)])
def test_every_view_of_a_batch_is_bit_exact_vs_reference(name, V, step, ref_mod):
    d = S.make_scene(name)
    check_batch_vs_reference(d, _cams(d, V, step), ref_mod)


def check_batch_vs_reference(d, cams, ref_mod):
    """Every view of one batched call against one reference call per view (also driven by tests/test_zy_fuzz_gpu.py)."""
    V = len(cams)
    W, H = cams[0].image_width, cams[0].image_height
    gc, gd, ga = _upstream(V, W, H)
    o = run_batch(d, cams, flags=_abi.FLAG_EXPORT_KEYS, grads=(gc, gd, ga))
    st = o["st"]
    total_R = 0
    ref_sum = None
    for v, cam in enumerate(cams):
        dv = dict(d); dv["cam"] = cam
        ref = ref_mod.RefCall(dv)
        rc, rr, rd, ra = ref.forward()
        total_R += ref.R
        assert torch.equal(o["radii"][v], rr), f"view {v}: radii differ"
        assert torch.equal(M.export_state(st, "tiles_touched", v), ref.state("tiles_touched"))
        vis = rr > 0
        for nm in ("depths", "means2D", "conic_opacity", "rgb") + (("cov3D",) if d["scales"] is not None else ()):
            if nm == "rgb" and d["colors_precomp"] is not None:
                continue
            a, b = M.export_state(st, nm, v)[vis], ref.state(nm)[vis]
            assert (_bits(a) != _bits(b)).sum().item() == 0, f"view {v}: {nm}"
        assert torch.equal(M.export_state(st, "keys_sorted", v), ref.state("keys_sorted")), f"view {v}: keys"
        assert torch.equal(M.export_state(st, "point_list", v), ref.state("point_list")), f"view {v}: ids"
        assert torch.equal(M.export_state(st, "ranges", v), ref.state("ranges")), f"view {v}: ranges"
        assert torch.equal(M.export_state(st, "n_contrib", v), ref.state("n_contrib"))
        assert torch.equal(_bits(M.export_state(st, "final_T", v)), _bits(ref.state("final_T")))
        for nm, a, b in (("color", o["color"][v], rc), ("depth", o["depth"][v], rd), ("alpha", o["alpha"][v], ra)):
            assert (_bits(a) != _bits(b)).sum().item() == 0, f"view {v}: {nm} not bit-identical"
        rg = ref.backward(gc[v], gd[v], ga[v])
        assert rel_err(o["grads"]["means2D"][v], rg["means2D"]) <= GRAD_TOL, f"view {v}: means2D"
        ref_sum = {k: t.double() for k, t in rg.items()} if ref_sum is None else {k: ref_sum[k] + rg[k].double() for k in rg}
    assert st.num_rendered == total_R
    checks = ["opacity", "means3D", "cov3D"] + (["sh"] if d["shs"] is not None else ["colors"])
    checks += ["scales", "rotations"] if d["scales"] is not None else []
    for nm in checks:
        e = rel_err(o["grads"][nm], ref_sum[nm])
        assert e <= GRAD_TOL, f"{nm}: relative error {e} vs the sum of the reference's per-view gradients"


@pytest.mark.parametrize("name,V", [("zju", 8), ("zju_sr", 4), ("ragged", 5), ("stress_small", 3)])
def test_batch_equals_separate_single_view_calls(name, V):
    d = S.make_scene(name)
    cams = _cams(d, V, 3) if name != "stress_small" else [d["cam"]] * V
    W, H = cams[0].image_width, cams[0].image_height
    gc, gd, ga = _upstream(V, W, H)
    o = run_batch(d, cams, grads=(gc, gd, ga))
    acc = None
    for v, cam in enumerate(cams):
        color, radii, depth, alpha, st = M.forward_raw(*_dev(d), _settings(d, cam))
        g = dict(zip(NAMES, M.backward_raw(st, radii, gc[v], gd[v], ga[v])))
        assert torch.equal(o["radii"][v], radii)
        for nm, a, b in (("color", o["color"][v], color), ("depth", o["depth"][v], depth), ("alpha", o["alpha"][v], alpha)):
            assert torch.equal(_bits(a), _bits(b)), (v, nm)
        assert rel_err(o["grads"]["means2D"][v], g["means2D"]) <= 1e-5
        acc = {k: t.double().clone() for k, t in g.items()} if acc is None else {k: acc[k] + g[k].double() for k in g}
    for nm in NAMES[1:]:
        if o["grads"][nm].numel():
            assert rel_err(o["grads"][nm], acc[nm]) <= 1e-5, nm


def test_batch_autograd_api_and_determinism():
    d = S.make_scene("small")
    V = 4
    cams = _cams(d, V, 5)
    rs = [_settings(d, c) for c in cams]
    t = {k: d[k].to(DEV).requires_grad_(True) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
    m2 = torch.zeros(V, d["means3D"].shape[0], 3, device=DEV, requires_grad=True)
    color, radii, depth, alpha = M.rasterize_batch(t["means3D"], m2, t["opacities"], rs, shs=t["shs"], scales=t["scales"],
                                                   rotations=t["rotations"])
    assert color.shape == (V, 3, 128, 128) and depth.shape == (V, 1, 128, 128) and radii.shape == (V, 2000)
    gc, gd, ga = _upstream(V, 128, 128)
    torch.autograd.backward([color, depth, alpha], [gc, gd, ga])
    o = run_batch(d, cams, grads=(gc, gd, ga))
    assert torch.equal(color.detach(), o["color"])
    assert rel_err(t["means3D"].grad, o["grads"]["means3D"]) <= 1e-5
    assert rel_err(t["shs"].grad, o["grads"]["sh"]) <= 1e-5
    assert rel_err(m2.grad, o["grads"]["means2D"]) <= 1e-5
    with pytest.raises(Exception, match="excatly one of either SHs"):
        M.rasterize_batch(t["means3D"], m2, t["opacities"], rs, scales=t["scales"], rotations=t["rotations"])
    # in-place edits of a render are allowed (three independent output tensors, as in the reference)
    c2, _, _, a2 = M.rasterize_batch(t["means3D"], m2, t["opacities"], rs, shs=t["shs"], scales=t["scales"], rotations=t["rotations"])
    c2.clamp_(0.0, 1.0)
    (c2.sum() + a2.sum()).backward()
    # deterministic mode: bit-identical from run to run, also for a batch (the view sum has a fixed order)
    runs = []
    for _ in range(3):
        ob = run_batch(d, cams, grads=(gc, gd, ga), bflags=_abi.FLAG_DETERMINISTIC)
        runs.append([ob["grads"][k].clone() for k in NAMES])
    for r in runs[1:]:
        for a, b in zip(runs[0], r):
            if a.numel():
                assert torch.equal(_bits(a), _bits(b))
    for a, k in zip(runs[0], NAMES):
        if a.numel():
            assert rel_err(a, o["grads"][k]) <= 1e-5, k


def test_single_view_outputs_can_be_edited_in_place():
    d = S.make_scene("small")
    rast = M.GaussianRasterizer(_settings(d, d["cam"]))
    t = {k: d[k].to(DEV).requires_grad_(True) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
    m2 = torch.zeros_like(t["means3D"], requires_grad=True)
    color, radii, depth, alpha = rast(means3D=t["means3D"], means2D=m2, shs=t["shs"], opacities=t["opacities"],
                                      scales=t["scales"], rotations=t["rotations"])
    color.clamp_(0.0, 1.0)                       # the reference returns independent tensors: this must work
    color.permute(1, 2, 0)[alpha[0] < 0.5] = 0
    (color.sum() + depth.sum()).backward()
    assert torch.isfinite(t["means3D"].grad).all()


@pytest.mark.parametrize("name,V", [("small", 1), ("small_cov", 1), ("small_color", 3), ("ragged", 4)])
def test_compiled_binding_equals_python_autograd_function(name, V):
    """The public API runs through moss_b200/_C.so (csrc/torch_binding.cpp); the Python autograd.Function over ctypes is
    the same glue.  Both must give identical forward bits and (deterministic accumulation aside) the same gradients,
    for single-view and batched calls, and free their state the same way."""
    from moss_b200 import rasterizer as R
    if R._load_binding() is None:
        pytest.skip("compiled binding not built")
    d = S.make_scene(name)
    cams = _cams(d, V, 5)
    rs = [_settings(d, c) for c in cams]
    W, H = cams[0].image_width, cams[0].image_height
    gc, gd, ga = _upstream(V, W, H)
    batched = V > 1
    if not batched:
        gc, gd, ga = gc[0], gd[0], ga[0]
    res = []
    for use_cpp in (True, False):
        t = {k: (d[k].to(DEV).requires_grad_(True) if d[k] is not None else R._EMPTY)
             for k in ("means3D", "shs", "colors_precomp", "opacities", "scales", "rotations", "cov3D_precomp")}
        P = d["means3D"].shape[0]
        m2 = torch.zeros(((V,) if batched else ()) + (P, 3), device=DEV, requires_grad=True)
        args = (t["means3D"], m2, t["shs"], t["colors_precomp"], t["opacities"], t["scales"], t["rotations"], t["cov3D_precomp"])
        if use_cpp:
            out = R._rasterize(*args, rs if batched else rs[0], batched)
        else:
            out = R._RasterizeGaussians.apply(*args, rs if batched else rs[0])
        color, radii, depth, alpha = out
        torch.autograd.backward([color, depth, alpha], [gc, gd, ga])
        res.append((color.detach(), radii, depth.detach(), alpha.detach(), m2.grad, {k: v.grad for k, v in t.items() if v.numel()}))
        with pytest.raises(RuntimeError):
            torch.autograd.backward([color], [gc])          # state was freed by the first backward
    a, b = res
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1]) and torch.equal(a[2], b[2]) and torch.equal(a[3], b[3])
    assert rel_err(a[4], b[4]) <= 1e-5
    for k in a[5]:
        assert rel_err(a[5][k], b[5][k]) <= 1e-5, k
