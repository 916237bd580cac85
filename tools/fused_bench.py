"""NS-a measurement: the fused skin + rasterize call against the two-call form (skin kernel -> HBM -> rasterizer) on the
same Gaussians, forward + backward through the public autograd wrappers.  Per-stage kernel times from the library's
event brackets, whole-step device time from CUDA events with a cold L2.  One JSON line per size.
    python tools/fused_bench.py [P ...]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import moss_b200 as M
from moss_b200 import fused as F, skinning as K, synth, _abi
from oracle import lbs_oracle as L
from tests import scenes as S
from tests.lbs_cases import oracle_smpl

dev = torch.device("cuda:0")
lib = _abi.load()
body = synth.body()
flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=dev)
for P, W, H in [(int(a), 512, 512) for a in sys.argv[1:]] or [(20000, 512, 512), (100000, 1024, 1024)]:
    g = synth.gaussians(P, body, seed=synth.SEED + 9)
    fr = synth.frame_params(9)
    fr.params["Th"] = torch.zeros(1, 3); fr.params["R"] = torch.eye(3)
    ft = K.FrameTables(*[t.to(dev).contiguous() for t in L.frame_tables(oracle_smpl(body), fr.params, fr.big_params)])
    cam = synth.ring_cameras(23, W, H)[2]
    rs = M.GaussianRasterizationSettings(**synth.settings_kwargs(cam, torch.zeros(3), 3, 1.0, device=dev))
    gc, gd, ga = (t.to(dev) for t in S.upstream_grads(W, H, depth=False))
    leaf = lambda t: t.to(dev).clone().requires_grad_(True)  # noqa: E731
    xyz, sc, rot, shs, opac = leaf(g.xyz), leaf(g.scales), leaf(g.rotations), leaf(g.shs), leaf(g.opacities)
    vid, wts = g.vert_ids.to(dev), body.weights.to(dev)
    rast = M.GaussianRasterizer(rs)

    def two_call():
        for t in (xyz, sc, rot, shs, opac): t.grad = None
        m, cov, _, _, _ = K.skin_and_covariance(xyz, vid, wts, ft, sc, rot, 1.0)
        c, r, d, a = rast(means3D=m, means2D=torch.zeros_like(m), shs=shs, colors_precomp=None, opacities=opac, scales=None,
                          rotations=None, cov3D_precomp=cov)
        c.backward(gc)

    def fused():
        for t in (xyz, sc, rot, shs, opac): t.grad = None
        c, r, d, a, _ = F.skin_and_rasterize(rs, xyz, vid, wts, ft, sc, rot, opac, shs=shs, return_transforms=False)
        c.backward(gc)

    def measure(step):
        for _ in range(10): step()
        torch.cuda.synchronize()
        l0 = lib.moss_kernel_launches()
        lib.moss_profile_enable(1)
        for _ in range(50):
            flush.zero_(); step()
        torch.cuda.synchronize()
        prof = _abi.profile_read(); lib.moss_profile_enable(0)
        launches = (lib.moss_kernel_launches() - l0) / 50
        us = {k: round(v[0] / v[1] * 1e3, 2) for k, v in prof.items() if v[1]}
        return us, launches
    u2, l2 = measure(two_call)
    u1, l1 = measure(fused)
    front2 = u2["skin_forward"] + u2["preprocess_forward"] + u2["preprocess_backward"] + u2["skin_backward"]
    front1 = u1["preprocess_forward"] + u1["preprocess_backward"]
    print(json.dumps({"metric": "fused skin + rasterize vs two calls (fwd+bwd)", "gaussians": P, "image": [W, H],
                      "two_call": {"launches": l2, "kernel_us": u2, "per_gaussian_kernels_us": round(front2, 2), "kernels_sum_us": round(sum(u2.values()), 2)},
                      "fused": {"launches": l1, "kernel_us": u1, "per_gaussian_kernels_us": round(front1, 2), "kernels_sum_us": round(sum(u1.values()), 2)},
                      "hbm_bytes_not_moved": P * 36 * 4, "per_gaussian_kernels_saved_us": round(front2 - front1, 2)}))
