"""Developer tool: which stage of a public-API training iteration cannot be captured in a CUDA graph?  Captures growing
prefixes of the iteration under rasterizer.async_launch() and reports the first one that fails."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import moss_b200 as M
from moss_b200 import fused as F, losses as ML, skinning as K, synth, rasterizer as R
from oracle import lbs_oracle as L
from tests.golden.make_loss_golden import synth_case

dev = torch.device("cuda:0")
body = synth.body()
V = body.verts.shape[0]
P, W, H = 20000, 512, 512
g = synth.gaussians(P, body, seed=synth.SEED + 9)
fr = synth.frame_params(9)
fr.params["Th"] = torch.zeros(1, 3); fr.params["R"] = torch.eye(3)
params = {k: v.to(dev) for k, v in fr.params.items()}
big = {k: v.to(dev) for k, v in fr.big_params.items()}
cam = synth.ring_cameras(23, W, H)[2]
rs = M.GaussianRasterizationSettings(**synth.settings_kwargs(cam, torch.zeros(3), 3, 1.0, device=dev))
box = (W // 4 + 1, H // 9, W - W // 4 - 2, H - H // 7 - 1)
_, _, gt, bkgd, bound = (torch.from_numpy(a).to(dev) for a in synth_case(W, H, 11, box))
verts, wts = body.verts.to(dev), body.weights.to(dev)
smpl = K.Smpl(verts, body.shapedirs.to(dev), body.posedirs.t().reshape(V, 3, 207).contiguous().to(dev), body.J_regressor.to(dev), wts, body.parents.to(dev))
leaf = lambda t: t.to(dev).clone().requires_grad_(True)  # noqa: E731
xyz, sc, rot, shs, opac = leaf(g.xyz), leaf(g.scales), leaf(g.rotations), leaf(g.shs), leaf(g.opacities)
poses = params["poses"].clone().requires_grad_(True)
cr = L.batch_rodrigues(0.05 * torch.randn(23, 3)).reshape(1, 23, 3, 3).to(dev).requires_grad_(True)
bounds = ML.mask_bounds(bound)
cache = K.FrameCache()
acc, den, mx = torch.zeros(P, 1, device=dev), torch.zeros(P, 1, device=dev), torch.zeros(P, device=dev)
NAMES = ["frame_tables", "nearest_vertex", "skin_and_rasterize", "image_losses", "backward", "densify_stats"]


def run(k):
    for t in (xyz, sc, rot, shs, opac, poses, cr):
        t.grad = None
    pr = dict(params); pr["poses"] = poses
    ft = K.frame_tables(smpl, pr, big, cr, cache=cache)
    if k < 2: return
    vid = K.nearest_vertex(xyz.detach(), verts)
    if k < 3: return
    m2 = torch.zeros_like(xyz, requires_grad=True)
    c, r, d, a, _ = F.skin_and_rasterize(rs, xyz, vid, wts, ft, sc, rot, opac, means2D=m2, shs=shs, return_transforms=False)
    if k < 4: return
    loss, _ = ML.image_losses(c, a, gt, bkgd, bound, bounds)
    if k < 5: return
    loss.backward()
    if k < 6: return
    M.add_densification_stats(r, m2.grad, acc, den, mx)


run(6); torch.cuda.synchronize()
for k in range(1, 7):
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side), R.async_launch():
        run(k); run(k)
    torch.cuda.current_stream().wait_stream(side); torch.cuda.synchronize()
    gph = torch.cuda.CUDAGraph()
    try:
        with R.async_launch(), torch.cuda.graph(gph):
            run(k)
        gph.replay(); torch.cuda.synchronize()
        print("prefix up to", NAMES[k - 1], ": captured and replayed", flush=True)
    except Exception as e:
        print("prefix up to", NAMES[k - 1], ": FAILED:", str(e).splitlines()[0], flush=True)
        break
