"""GPU: randomised scenes against the UNMODIFIED reference rasterizer (the reference has no tests,
so the new code brings property tests over random and degenerate inputs).  hypothesis draws the scene description --
Gaussian count, image size (any W, H: not multiples of 16), active SH degree and stored coefficient count, SH or precomputed
colours, scale + rotation or precomputed covariance, splat size from sub-pixel to screen-filling, camera inside / near / far,
Gaussians behind the camera (up to all of them: nothing rendered), exactly transparent and exactly opaque splats, background
-- and every example must give the reference's bits for radii, tile counts, the sorted key and ID lists, tile ranges,
n_contrib, final_T and the three images, and its gradients within 1e-3 relative.  derandomize=True: the same examples in
every run, so a pass here is a pass at the end of the round.  (Named test_zy_* so that it is collected behind the other
suites: the reference is driven with inputs nobody ran it on before, and a fault inside it would poison the CUDA context for
every test after it.)"""
import os

import pytest
import torch

hypothesis = pytest.importorskip("hypothesis")
from hypothesis import given, settings, strategies as st, HealthCheck  # noqa: E402

import moss_b200 as M  # noqa: E402
from tests import scenes as S  # noqa: E402
from moss_b200 import synth  # noqa: E402
from tests.test_parity_gpu import run_ours, rel_err, _bits, GRAD_TOL, ref_mod  # noqa: E402,F401
from tests.test_batch_gpu import check_batch_vs_reference  # noqa: E402

pytestmark = pytest.mark.gpu


@st.composite
def scene_specs(draw):
    D = draw(st.integers(0, 3))
    M_store = draw(st.sampled_from([m for m in (1, 4, 9, 16) if m >= (D + 1) ** 2]))
    spec = dict(
        P=draw(st.integers(4, 1500)), W=draw(st.integers(16, 260)), H=draw(st.integers(16, 260)),
        D=D, sh_M=M_store, seed=draw(st.integers(0, 5)), cam=draw(st.integers(0, 22)),
        radius=draw(st.sampled_from([0.6, 1.5, 3.0, 8.0])), mod=draw(st.sampled_from([0.5, 1.0, 2.0])),
        scale_mul=draw(st.sampled_from([0.02, 0.3, 1.0, 4.0, 25.0])),
        precomp_color=draw(st.booleans()), precomp_cov=draw(st.booleans()), transforms=draw(st.booleans()),
        raw_quat=draw(st.booleans()), behind=draw(st.sampled_from([0, 0.1, 0.5, 1.0])),
        opacity_edges=draw(st.booleans()),
        bg=[draw(st.sampled_from([0.0, 0.25, 1.0])) for _ in range(3)],
    )
    return spec


# MOSS_FUZZ_EXAMPLES=n / MOSS_FUZZ_RANDOM=1: a longer or a freshly seeded search (developer runs; the defaults are what CI runs)
@settings(max_examples=int(os.environ.get("MOSS_FUZZ_EXAMPLES", "60")), deadline=None,
          derandomize=not os.environ.get("MOSS_FUZZ_RANDOM"), database=None, suppress_health_check=list(HealthCheck))
@given(spec=scene_specs(), with_depth=st.booleans())
def test_random_scenes_equal_the_reference(ref_mod, spec, with_depth):  # noqa: F811
    d = S.make_scene_from_spec(spec)
    assert all(torch.isfinite(t).all() for t in d.values() if torch.is_tensor(t))     # NaN inputs: see test_nan_covariance_is_culled
    cam = d["cam"]
    grads = S.upstream_grads(cam.image_width, cam.image_height, seed=11 + spec["P"], depth=with_depth)
    ref = ref_mod.RefCall(d)
    rc, rr, rd, ra = ref.forward()
    o = run_ours(d, grads=grads)
    st_ = o["st"]
    what = f"spec {spec}"
    assert torch.equal(o["radii"], rr), what
    assert torch.equal(M.export_state(st_, "tiles_touched"), ref.state("tiles_touched")), what
    assert o["R"] == ref.R, what
    for nm in ("keys_sorted", "point_list", "ranges", "n_contrib"):
        assert torch.equal(M.export_state(st_, nm), ref.state(nm)), (nm, what)
    assert torch.equal(_bits(M.export_state(st_, "final_T")), _bits(ref.state("final_T"))), what
    for nm, a, b in (("color", o["color"], rc), ("depth", o["depth"], rd), ("alpha", o["alpha"], ra)):
        assert torch.equal(_bits(a), _bits(b)), (nm, float((a - b).abs().max()), what)
    rg = ref.backward(*grads)
    og = o["grads"]
    checks = ["means2D", "opacity", "means3D", "cov3D"]
    checks += ["sh"] if d["shs"] is not None else ["colors"]
    checks += ["scales", "rotations"] if d["scales"] is not None else []
    for nm in checks:
        assert torch.isfinite(og[nm]).all() == torch.isfinite(rg[nm]).all(), (nm, what)
        if torch.isfinite(rg[nm]).all():
            # tiny gradients (a handful of faint pixels) are compared against the tensor's own rounding floor
            e = rel_err(og[nm], rg[nm])
            floor = 1e-12 / max(float(rg[nm].double().norm()), 1e-30)
            assert e <= GRAD_TOL + floor, (nm, e, what)
    inv = ~(o["radii"] > 0)
    for nm in checks:
        assert float(og[nm][inv].abs().sum()) == 0.0, (nm, what)
    # the same scene through the public drop-in API (compiled autograd binding): the raw calls' outputs, bit for bit
    from tests.test_parity_gpu import _settings
    leaf = lambda t: None if t is None else t.to("cuda").requires_grad_(True)  # noqa: E731
    t = {k: leaf(d[k]) for k in ("means3D", "shs", "colors_precomp", "opacities", "scales", "rotations", "cov3D_precomp")}
    m2d = torch.zeros_like(t["means3D"], requires_grad=True)
    color, radii, depth, alpha = M.GaussianRasterizer(_settings(d, "cuda"))(
        means3D=t["means3D"], means2D=m2d, shs=t["shs"], colors_precomp=t["colors_precomp"], opacities=t["opacities"],
        scales=t["scales"], rotations=t["rotations"], cov3D_precomp=t["cov3D_precomp"])
    assert torch.equal(radii, o["radii"]), what
    for a, b in ((color, o["color"]), (depth, o["depth"]), (alpha, o["alpha"])):
        assert torch.equal(a, b), what
    torch.autograd.backward([color, depth, alpha], [g.to("cuda") for g in grads])
    api = dict(means2D=m2d.grad, opacity=t["opacities"].grad, means3D=t["means3D"].grad,
               cov3D=None if t["cov3D_precomp"] is None else t["cov3D_precomp"].grad,
               sh=None if t["shs"] is None else t["shs"].grad,
               colors=None if t["colors_precomp"] is None else t["colors_precomp"].grad,
               scales=None if t["scales"] is None else t["scales"].grad,
               rotations=None if t["rotations"] is None else t["rotations"].grad)
    for nm, g in api.items():
        if g is None or (nm == "cov3D" and d["cov3D_precomp"] is None):
            continue
        ref_g = og[nm].reshape(g.shape)
        assert rel_err(g, ref_g) <= 1e-5 + 1e-12 / max(float(ref_g.double().norm()), 1e-30), (nm, what)


@settings(max_examples=int(os.environ.get("MOSS_FUZZ_EXAMPLES", "25")), deadline=None,
          derandomize=not os.environ.get("MOSS_FUZZ_RANDOM"), database=None, suppress_health_check=list(HealthCheck))
@given(spec=scene_specs(), V=st.integers(2, 5), first=st.integers(0, 22), step=st.integers(1, 11))
def test_random_batches_equal_the_reference_view_by_view(ref_mod, spec, V, first, step):  # noqa: F811
    """The batched call (view index folded into tile and Gaussian indices) on random scenes and random camera subsets."""
    d = S.make_scene_from_spec(spec)
    ring = synth.ring_cameras(23, spec["W"], spec["H"], radius=spec["radius"])
    cams = [ring[(first + k * step) % 23] for k in range(V)]
    check_batch_vs_reference(d, cams, ref_mod)


def test_nan_covariance_is_culled():
    """A NaN covariance gives radius 0 (float -> int of NaN).  The reference still counts the Gaussian in num_rendered but
    never emits its key (DGR rasterizer_impl.cu:82), so its sorted list keeps an uninitialised entry that the blend kernels
    use as an index; this library culls it: nothing rendered for it, finite outputs, zero gradient rows."""
    d = S.make_scene("small")
    d["scales"][::7] = float("nan")
    grads = S.upstream_grads(d["cam"].image_width, d["cam"].image_height)
    o = run_ours(d, grads=grads)
    bad = torch.zeros(d["means3D"].shape[0], dtype=torch.bool, device=o["radii"].device)
    bad[::7] = True
    assert int((o["radii"][bad] != 0).sum()) == 0
    assert int(M.export_state(o["st"], "tiles_touched")[bad].sum()) == 0
    assert int((M.export_state(o["st"], "point_list").long() % 7 == 0).sum()) == 0      # none of them in the sorted list
    for nm in ("color", "depth", "alpha"):
        assert torch.isfinite(o[nm]).all(), nm
    for nm, g in o["grads"].items():
        if g is not None and g.numel():
            assert torch.isfinite(g).all(), nm
            assert float(g[bad].abs().sum()) == 0.0, nm
    clean = S.make_scene("small")
    keep = ~bad.cpu()
    for k in ("means3D", "shs", "opacities", "scales", "rotations"):
        clean[k] = clean[k][keep].contiguous()
    oc = run_ours(clean)
    assert torch.equal(o["color"], oc["color"]) and torch.equal(o["alpha"], oc["alpha"])   # as if they were not there


@st.composite
def loss_cases(draw):
    W, H = draw(st.integers(13, 200)), draw(st.integers(13, 200))
    x0, y0 = draw(st.integers(0, W - 2)), draw(st.integers(0, H - 2))
    x1, y1 = draw(st.integers(x0 + 1, W - 1)), draw(st.integers(y0 + 1, H - 1))
    weights = draw(st.sampled_from([(1.0, 0.5, 0.2, 0.3), (1.0, 0.5, 0.2, 0.3), (0.0, 0.0, 1.0, 0.0), (0.0, 0.0, 0.0, 1.0), (1.0, 0.0, 0.0, 0.0)]))
    return W, H, (x0, y0, x1, y1), draw(st.integers(0, 1000)), weights


@settings(max_examples=int(os.environ.get("MOSS_FUZZ_EXAMPLES", "40")), deadline=None,
          derandomize=not os.environ.get("MOSS_FUZZ_RANDOM"), database=None, suppress_health_check=list(HealthCheck))
@given(case=loss_cases())
def test_random_images_and_crops_equal_the_loss_oracle(case):
    """The fused image-loss kernel on random image sizes and crop rectangles (anywhere in the image, touching its borders,
    narrower than the 11-tap window, not aligned to the kernel's 32-pixel tiles) against the CPU oracle, which is pinned to
    the real utils/loss_utils.py by tests/golden/loss_*.npz."""
    from moss_b200 import losses as ML
    from oracle import loss_oracle as LO
    from tests.golden.make_loss_golden import synth_case
    W, H, box, seed, w = case
    image, alpha, gt, bkgd, bound = (torch.from_numpy(a) for a in synth_case(W, H, seed, box))
    ic, ac = image.clone().requires_grad_(True), alpha.clone().requires_grad_(True)
    total, t = LO.image_losses(ic, ac, gt, bkgd, bound, weights=w)
    gi, ga = torch.autograd.grad(total, [ic, ac], allow_unused=True)
    dev = "cuda:0"
    ig, ag = image.to(dev).requires_grad_(True), alpha.to(dev).requires_grad_(True)
    b = ML.mask_bounds(bound.to(dev))
    assert b[:4].tolist() == [box[0], box[1], box[2] - box[0] + 1, box[3] - box[1] + 1], (case, b.tolist())
    loss, terms = ML.image_losses(ig, ag, gt.to(dev), bkgd.to(dev), bound.to(dev), b, w)
    loss.backward()
    assert abs(float(loss.detach()) - float(total.detach())) <= 1e-5 * max(1.0, abs(float(total.detach()))), case
    for got, ref in ((ig.grad, gi), (ag.grad, ga)):
        ref = torch.zeros_like(got.cpu()) if ref is None else ref
        scale = max(float(ref.abs().max()), 1e-30)
        err = float((got.cpu().double() - ref.double().reshape(got.shape)).abs().max())
        # absolute floor: a crop inside the rows where image == gt has a zero gradient, and both sides hold 1e-8 .. 1e-7 of rounding
        assert err <= 5e-4 * scale + 1e-6, (case, err, scale)


@settings(max_examples=int(os.environ.get("MOSS_FUZZ_EXAMPLES", "12")), deadline=None,
          derandomize=not os.environ.get("MOSS_FUZZ_RANDOM"), database=None, suppress_health_check=list(HealthCheck))
@given(P=st.integers(50, 3000), W=st.integers(48, 300), aspect=st.sampled_from([0.7, 1.0, 1.3]), seed=st.integers(0, 40),
       with_delta=st.booleans())
def test_random_fused_skin_rasterize_calls(P, W, aspect, seed, with_delta):
    """The fused skin + rasterize call on random sizes, poses and LBS offsets: the two-step parity of tests/test_fused_gpu.py
    (registers == front-end kernel, images / radii / binning == the unmodified reference rasterizer, gradients == the
    two-call form)."""
    from tests import test_fused_gpu as TF
    H = max(int(W * aspect), 32)
    TF.test_forward_two_step_parity(P, W, H, seed, with_delta)
    TF.test_backward_matches_two_call_form(P, W, H, seed, with_delta)
