"""Developer tool for `compute-sanitizer`: one fwd+bwd of the named scenes through the raw ABI, a
3-view batch, and the front-end kernels (frame tables, skinning, nearest vertex, densify stats, AdamW) -- every kernel
of the library runs at least once, at a Stress-size and a ragged-size problem.
    compute-sanitizer --tool memcheck  python tools/sanitize_step.py stress ragged
    compute-sanitizer --tool racecheck python tools/sanitize_step.py stress_small ragged"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import moss_b200 as M
from moss_b200 import synth, skinning as K, _abi
from tests import scenes as S

dev = torch.device("cuda:0")
names = sys.argv[1:] or ["ragged"]
for name in names:
    d = S.make_scene(name)
    cam = d["cam"]
    rs = M.GaussianRasterizationSettings(**synth.settings_kwargs(cam, d["bg"], d["sh_degree"], d["scale_modifier"], device=dev))
    t = {k: (d[k].to(dev) if d[k] is not None else torch.empty(0)) for k in ("means3D", "shs", "colors_precomp", "opacities", "scales", "rotations", "cov3D_precomp")}
    gc, gd, ga = (x.to(dev) for x in S.upstream_grads(cam.image_width, cam.image_height, depth=True))
    for flags in (0, _abi.FLAG_DETERMINISTIC):
        c, r, dp, a, st = M.forward_raw(t["means3D"], t["shs"], t["colors_precomp"], t["opacities"], t["scales"], t["rotations"], t["cov3D_precomp"], rs, flags=flags)
        M.backward_raw(st, r, gc, gd, ga)
    torch.cuda.synchronize()
    print("raster ok", name, int((r > 0).sum()), flush=True)

# a 3-view batch through the public API (batched forward/backward, view sum)
d = S.make_scene("ragged")
cams = synth.ring_cameras(23, d["cam"].image_width, d["cam"].image_height)[:3]
settings = [M.GaussianRasterizationSettings(**synth.settings_kwargs(c, d["bg"], d["sh_degree"], d["scale_modifier"], device=dev)) for c in cams]
leaf = lambda x: None if x is None else x.to(dev).requires_grad_(True)  # noqa: E731
m3, shs, op, sc, rot = (leaf(d[k]) for k in ("means3D", "shs", "opacities", "scales", "rotations"))
out = M.rasterize_batch(m3, None, op, settings, shs=shs, colors_precomp=leaf(d["colors_precomp"]), scales=sc, rotations=rot,
                        cov3D_precomp=leaf(d["cov3D_precomp"]))
out[0].sum().backward()
torch.cuda.synchronize()
print("batch ok", flush=True)

# front end: frame tables -> skinning -> gradients down to the pose parameters; nearest vertex
body = synth.body()
V = body.verts.shape[0]
smpl = K.Smpl(body.verts.to(dev), body.shapedirs.to(dev), body.posedirs.t().reshape(V, 3, 207).contiguous().to(dev),
              body.J_regressor.to(dev), body.weights.to(dev), body.parents.to(dev))
fr = synth.frame_params(1)
params = {k: v.to(dev) for k, v in fr.params.items()}
big = {k: v.to(dev) for k, v in fr.big_params.items()}
params["poses"].requires_grad_(True)
g = synth.gaussians(5001, body)
ft = K.frame_tables(smpl, params, big)
xyz = g.xyz.to(dev).requires_grad_(True)
vid = K.nearest_vertex(xyz, smpl.v_template)
m, cov, _, _, _ = K.skin_and_covariance(xyz, vid, smpl.weights, ft, g.scales.to(dev).requires_grad_(True), g.rotations.to(dev).requires_grad_(True))
(m.sum() + cov.sum()).backward()
torch.cuda.synchronize()
print("front end ok", float(params["poses"].grad.abs().sum()), flush=True)
