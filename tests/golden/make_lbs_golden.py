"""Generate tests/golden/lbs_*.npz by running the REAL reference skinning/covariance code on the CPU.

Run in the build container only (needs /root/reference; the GPU box never reads it):

    python tests/golden/make_lbs_golden.py

What is imported from /root/reference, unmodified: scene/gaussian_model.py (GaussianModel.coarse_deform_c2source
:820-923, get_transform_params_torch :998-1031, get_rigid_transformation_torch :965-995, batch_rodrigues :1033-1064,
the covariance closure :37-44) and utils/general_utils.py (build_rotation, build_scaling_rotation, strip_symmetric).
What is stubbed, because it is absent from this image: open3d, cv2, plyfile, simple_knn, knn_cuda (the k=1 nearest
vertex search is replaced by an exact cdist argmin), pytorch3d; `.cuda()` and `device="cuda"` are turned into
no-ops so the same python arithmetic runs in fp32 on the host.
"""
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF = "/root/reference"
sys.path.insert(0, ROOT)


def _stub_modules():
    def mod(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules[name] = m
        return m

    class KNN:
        def __init__(self, k=1, transpose_mode=True):
            self.k = k

        def __call__(self, ref, query):
            d = torch.cdist(query.double(), ref.double())
            dist, idx = d.topk(self.k, dim=-1, largest=False)
            return dist.float(), idx

    for name in ("open3d", "cv2"):
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:
                mod(name)
    mod("plyfile", PlyData=object, PlyElement=object)
    mod("simple_knn")
    mod("simple_knn._C", distCUDA2=None)
    mod("knn_cuda", KNN=KNN)
    mod("pytorch3d")
    mod("pytorch3d.transforms", matrix_to_quaternion=None)
    # host execution of code that hard-codes the device
    torch.Tensor.cuda = lambda self, *a, **k: self
    for fn in ("zeros", "eye", "ones", "empty", "tensor"):
        orig = getattr(torch, fn)

        def wrapped(*a, __orig=orig, **k):
            if k.get("device") in ("cuda", torch.device("cuda")):
                k.pop("device")
            return __orig(*a, **k)
        setattr(torch, fn, wrapped)


def reference_model(body):
    _stub_modules()
    sys.path.insert(0, REF)
    import importlib.util
    # scene/__init__.py drags in the dataset readers (imageio, cv2 ...): load the one module file directly
    spec = importlib.util.spec_from_file_location("ref_gaussian_model", os.path.join(REF, "scene", "gaussian_model.py"))
    GM = importlib.util.module_from_spec(spec)      # the real reference module
    spec.loader.exec_module(GM)
    m = GM.GaussianModel.__new__(GM.GaussianModel)
    m.setup_functions()
    m.knn = sys.modules["knn_cuda"].KNN(k=1)
    V = body.verts.shape[0]
    m.SMPL_NEUTRAL = {
        "v_template": body.verts, "shapedirs": body.shapedirs,
        "posedirs": body.posedirs.t().reshape(V, 3, 207).contiguous(),    # reference layout [V,3,207]
        "J_regressor": body.J_regressor, "weights": body.weights,
        "kintree_table": torch.stack([body.parents, torch.arange(24)]).numpy(),
    }
    return GM, m


def make_case(name, P, seed, with_delta, with_correct_Rs, mod):
    from moss_b200 import synth
    body = synth.body()
    GM, m = reference_model(body)
    fr = synth.frame_params(seed)
    g = synth.gaussians(P, body, seed=synth.SEED + seed)
    rng = np.random.default_rng(99 + seed)
    lbs_delta = torch.from_numpy(0.3 * rng.standard_normal((1, P, 24))).float() if with_delta else None
    correct_Rs = None
    if with_correct_Rs:
        aa = torch.from_numpy(0.05 * rng.standard_normal((23, 3))).float()
        correct_Rs = GM.batch_rodrigues(aa).reshape(1, 23, 3, 3)
    # the big-pose world vertices the reference looks Gaussians up against (scene/dataset_readers.py:586-601):
    # here simply the rest template, which is what xyz was sampled on.
    t_vertices = body.verts[None]
    raw_rot = g.rotations * torch.from_numpy(rng.uniform(0.7, 1.4, (P, 1))).float()   # stored quaternions are not unit
    xyz = g.xyz.clone().requires_grad_(True)
    scal = g.scales.clone().requires_grad_(True)
    rot = raw_rot.clone().requires_grad_(True)
    if lbs_delta is not None:
        lbs_delta.requires_grad_(True)
    m._xyz = xyz
    m._scaling = torch.log(scal)
    m._rotation = rot
    smpl_pts, world, bw, transforms, transl = m.coarse_deform_c2source(
        xyz[None], fr.params, fr.big_params, t_vertices, lbs_weights=lbs_delta, correct_Rs=correct_Rs,
        return_transl=True)
    cov = m.get_covariance(mod, transforms.squeeze(0))
    # upstream gradients as the rasterizer would hand them back (dL/dmeans3D, dL/dcov3D), seeded
    g_world = torch.from_numpy(rng.standard_normal((P, 3))).float()
    g_cov = torch.from_numpy(rng.standard_normal((P, 6)) * 50.0).float()
    leaves = [xyz, scal, rot] + ([lbs_delta] if lbs_delta is not None else [])
    grads = torch.autograd.grad([world[0], cov], leaves, [g_world, g_cov])
    with torch.no_grad():
        _, vert_ids = m.knn(t_vertices.float(), g.xyz[None].float())
    smpl_pts, world, bw, transforms, transl, cov = (t.detach() for t in (smpl_pts, world, bw, transforms, transl, cov))
    out = dict(P=P, seed=seed, mod=np.float32(mod), xyz=g.xyz.numpy(), scaling=g.scales.numpy(), rotation=raw_rot.numpy(),
               vert_ids=vert_ids.reshape(-1).numpy().astype(np.int32),
               smpl_pts=smpl_pts[0].numpy(), world=world[0].numpy(), bweights=bw[0].numpy(),
               transforms=transforms[0].numpy(), translation=transl[0].numpy(), cov3D=cov.numpy())
    out.update(g_world=g_world.numpy(), g_cov=g_cov.numpy(), d_xyz=grads[0].numpy(), d_scaling=grads[1].numpy(),
               d_rotation=grads[2].numpy())
    if lbs_delta is not None:
        out["lbs_delta"] = lbs_delta[0].detach().numpy()
        out["d_lbs_delta"] = grads[3][0].numpy()
    if correct_Rs is not None:
        out["correct_Rs"] = correct_Rs[0].numpy()
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), f"lbs_{name}.npz")
    np.savez_compressed(path, **out)
    print(name, "->", path, {k: getattr(v, "shape", v) for k, v in out.items()})


if __name__ == "__main__":
    make_case("plain", 1500, 0, False, False, 1.0)
    make_case("delta", 1200, 1, True, False, 1.0)
    make_case("correct", 1000, 2, True, True, 0.8)
