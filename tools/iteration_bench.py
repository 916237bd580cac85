"""One MOSS training iteration as ONE timed loop (DESIGN.md section 9 adds up separately measured parts; this runs them
back to back): per-frame SMPL tables -> nearest SMPL vertex -> skinning + covariance -> rasterizer -> masked L1 / L2 + SSIM
+ S3IM -> backward through all of it -> densification statistics -> AdamW over the six per-Gaussian groups
(train_ZJU.py:90-189, gaussian_renderer/__init__.py render()).  LPIPS, the pose / LBS-offset MLPs and the pose NLL are
left out on both sides.
  ours      = this library's public entry points (skinning.frame_tables, nearest_vertex, fused.skin_and_rasterize,
              losses.image_losses, add_densification_stats, FusedAdamW);
  reference = the reference's own ops on the same GPU: its python code restated line by line in oracle/ (pinned to the real
              functions by the golden tests) as torch kernels with autograd, the UNMODIFIED CUDA rasterizer from oracle/_ref,
              torch.cdist + argmin standing in for the knn_cuda wheel, torch.optim.AdamW.
Reported per arm: wall-clock ms per iteration (host included: a loop of N iterations, one synchronize at the end) and the
device time between CUDA events around the same loop.  One JSON line per size.
    python tools/iteration_bench.py [P ...]"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import moss_b200 as M
from moss_b200 import fused as F, losses as ML, skinning as K, synth, _abi
from moss_b200.optim import FusedAdamW
from oracle import lbs_oracle as L, loss_oracle as LO
from tests.lbs_cases import oracle_smpl
from tests.golden.make_loss_golden import synth_case

dev = torch.device("cuda:0")
lib = _abi.load()
body = synth.body()
V = body.verts.shape[0]
# MOSS's learning-rate ratios (arguments/__init__.py) scaled down: the targets here are synthetic, and hundreds of full-size
# steps towards them would blow the Gaussians up (the instance count tripled) and make the timings depend on the step count
LRS = {k: v * 1e-3 for k, v in {"xyz": 8e-4, "f_dc": 2.5e-3, "f_rest": 1.25e-4, "opacity": 0.05, "scaling": 5e-3, "rotation": 1e-3}.items()}
ref_dir = os.path.join(ROOT, "oracle", "_ref")
have_ref = os.path.exists(os.path.join(ref_dir, "diff_gaussian_rasterization", "_C.so"))
if have_ref:
    sys.path.insert(0, ref_dir)
    import diff_gaussian_rasterization as DGR


def make_params(g):
    shs = g.shs.to(dev)
    p = {"xyz": g.xyz.to(dev), "f_dc": shs[:, :1].contiguous(), "f_rest": shs[:, 1:].contiguous(), "opacity": g.opacities.to(dev),
         "scaling": g.scales.to(dev), "rotation": g.rotations.to(dev)}
    return {k: torch.nn.Parameter(v.clone()) for k, v in p.items()}


def loop_time(step, n, warmup=10):
    for _ in range(warmup):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for _ in range(n):
        step()
    e1.record()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3, e0.elapsed_time(e1) / n


for P, W, H in [(int(a), 512, 512) for a in sys.argv[1:]] or [(20000, 512, 512), (100000, 1024, 1024)]:
    g = synth.gaussians(P, body, seed=synth.SEED + 9)
    fr = synth.frame_params(9)
    fr.params["Th"] = torch.zeros(1, 3); fr.params["R"] = torch.eye(3)
    params = {k: v.to(dev) for k, v in fr.params.items()}
    big = {k: v.to(dev) for k, v in fr.big_params.items()}
    cam = synth.ring_cameras(23, W, H)[2]
    kw = synth.settings_kwargs(cam, torch.zeros(3), 3, 1.0, device=dev)
    box = (W // 4 + 1, H // 9, W - W // 4 - 2, H - H // 7 - 1)
    _, _, gt, bkgd, bound = (torch.from_numpy(a).to(dev) for a in synth_case(W, H, 11, box))
    verts, wts = body.verts.to(dev), body.weights.to(dev)
    out = {"metric": "one MOSS training iteration, ms", "gaussians": P, "image": [W, H], "vertices": V,
           "stages": "frame tables, nearest vertex, skin + covariance, rasterize, L1/L2/SSIM/S3IM, backward, densification stats, AdamW"}

    # ---- ours ------------------------------------------------------------------------------------------------------------
    smpl = K.Smpl(verts, body.shapedirs.to(dev), body.posedirs.t().reshape(V, 3, 207).contiguous().to(dev),
                  body.J_regressor.to(dev), wts, body.parents.to(dev))
    po = make_params(g)
    opt_o = FusedAdamW([{"params": [po[k]], "lr": LRS[k], "name": k} for k in LRS], lr=0.0, eps=1e-15)
    poses_o = params["poses"].clone().requires_grad_(True)
    cr_o = L.batch_rodrigues(0.05 * torch.randn(23, 3)).reshape(1, 23, 3, 3).to(dev).requires_grad_(True)
    rs_o = M.GaussianRasterizationSettings(**kw)
    bounds = ML.mask_bounds(bound)
    cache = K.FrameCache()
    acc, den, mx = torch.zeros(P, 1, device=dev), torch.zeros(P, 1, device=dev), torch.zeros(P, device=dev)
    last = {}

    def ours(step=True):
        opt_o.zero_grad(set_to_none=True); poses_o.grad = None; cr_o.grad = None
        pr = dict(params); pr["poses"] = poses_o
        ft = K.frame_tables(smpl, pr, big, cr_o, cache=cache)
        vid = K.nearest_vertex(po["xyz"].detach(), verts)
        means2D = torch.zeros_like(po["xyz"], requires_grad=True)
        shs = torch.cat((po["f_dc"], po["f_rest"]), dim=1)
        c, r, d, a, _ = F.skin_and_rasterize(rs_o, po["xyz"], vid, wts, ft, po["scaling"], po["rotation"], po["opacity"],
                                            means2D=means2D, shs=shs, return_transforms=False)
        loss, _ = ML.image_losses(c, a, gt, bkgd, bound, bounds)
        loss.backward()
        M.add_densification_stats(r, means2D.grad, acc, den, mx)
        if step:
            opt_o.step()
        last["ours"] = loss.detach()      # never keep an autograd graph of an eager run alive across a capture: its
        return last["ours"]               # AccumulateGrad nodes remember the (legacy) stream they were made on

    ours(); first_o = float(last["ours"])            # same start on both sides: the first losses must agree
    if os.environ.get("ITER_PROFILE"):               # where the host time of an iteration goes (launches are asynchronous)
        import cProfile, pstats
        for _ in range(20): ours()
        torch.cuda.synchronize()
        prof = cProfile.Profile(); prof.enable()
        for _ in range(200): ours()
        prof.disable(); torch.cuda.synchronize()
        st = pstats.Stats(prof); st.sort_stats("cumulative").print_stats(45); st.sort_stats("tottime").print_stats(30)
    l0 = lib.moss_kernel_launches()
    wall, devt = loop_time(ours, 200)
    out["ours"] = {"wall_ms": round(wall, 4), "device_ms": round(devt, 4), "library_launches_per_iteration": (lib.moss_kernel_launches() - l0) / 210,
                   "first_loss": round(first_o, 6)}

    # per-stage kernel time of one eager iteration (the library's event brackets)
    lib.moss_profile_enable(1)
    for _ in range(20): ours()
    torch.cuda.synchronize()
    prof = _abi.profile_read(); lib.moss_profile_enable(0)
    out["ours"]["kernel_us"] = {k: round(v[0] / v[1] * 1e3 * (v[1] / 20), 1) for k, v in prof.items() if v[1]}
    out["ours"]["kernel_sum_us"] = round(sum(out["ours"]["kernel_us"].values()), 1)
    # eager, without the host wait for the instance count in the middle of the forward (rasterizer.async_launch)
    from moss_b200 import rasterizer as R
    with R.async_launch():
        wall_a, devt_a = loop_time(ours, 200)
    cnt = [c.cpu().tolist() for c in R.async_counters()[-3:]]
    out["ours_async_eager"] = {"wall_ms": round(wall_a, 4), "device_ms": round(devt_a, 4), "last_counters": cnt}

    # ---- ours, the same iteration captured as one CUDA graph (the optimizer's step count lives on the host: it steps outside)
    from moss_b200.graphs import GraphedIteration
    gi = GraphedIteration(lambda: ours(step=False), dev)

    def graphed():
        gi.replay()
        opt_o.step()
    wall_g, devt_g = loop_time(graphed, 200)
    out["ours_graphed"] = {"wall_ms": round(wall_g, 4), "device_ms": round(devt_g, 4), "instances": gi.check(),
                           "note": "forward + losses + backward + densification statistics replayed as one graph, FusedAdamW stepped behind it"}

    # ---- reference ops on the same GPU -----------------------------------------------------------------------------------
    if have_ref:
        osm = L.Smpl(*[t.to(dev) for t in oracle_smpl(body)])
        pr_ = make_params(g)
        opt_r = torch.optim.AdamW([{"params": [pr_[k]], "lr": LRS[k], "name": k} for k in LRS], lr=0.0, eps=1e-15)
        poses_r = params["poses"].clone().requires_grad_(True)
        cr_r = cr_o.detach().clone().requires_grad_(True)
        rast = DGR.GaussianRasterizer(DGR.GaussianRasterizationSettings(**kw))
        acc_r, den_r, mx_r = torch.zeros(P, 1, device=dev), torch.zeros(P, 1, device=dev), torch.zeros(P, device=dev)

        def ref():
            opt_r.zero_grad(set_to_none=True); poses_r.grad = None; cr_r.grad = None
            pr = dict(params); pr["poses"] = poses_r
            ft = L.frame_tables(osm, pr, big, cr_r)
            vid = torch.cdist(pr_["xyz"].detach(), verts).argmin(dim=1)
            sk, cov = L.skin_and_covariance(pr_["xyz"], vid, wts, ft, pr_["scaling"], pr_["rotation"], 1.0, None)
            m = sk.world_pts
            means2D = torch.zeros_like(m, requires_grad=True)
            shs = torch.cat((pr_["f_dc"], pr_["f_rest"]), dim=1)
            c, r, d, a = rast(means3D=m, means2D=means2D, shs=shs, colors_precomp=None, opacities=pr_["opacity"], scales=None,
                              rotations=None, cov3D_precomp=cov)
            loss, _ = LO.image_losses(c, a, gt, bkgd, bound)
            loss.backward()
            vis = r > 0
            mx_r[vis] = torch.max(mx_r[vis], r[vis])
            acc_r[vis] += torch.norm(means2D.grad[vis, :2], dim=-1, keepdim=True)
            den_r[vis] += 1
            opt_r.step()
            last["ref"] = loss.detach()

        ref(); first_r = float(last["ref"])
        wall_r, devt_r = loop_time(ref, 30, warmup=5)
        out["reference_ops_same_gpu"] = {"wall_ms": round(wall_r, 4), "device_ms": round(devt_r, 4), "first_loss": round(first_r, 6)}
        out["speedup_wall"] = round(wall_r / wall, 1)
        out["speedup_device"] = round(devt_r / devt, 1)
        out["speedup_wall_graphed"] = round(wall_r / wall_g, 1)
    print(json.dumps(out), flush=True)
