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
    # poses / correct_Rs are leaves too: the real reference's autograd then yields the gradient that reaches them through
    # the kinematic chain and the pose blend shapes (pins d_A_tgt / d_poseoff_tgt and the frame kernel end to end)
    poses = fr.params["poses"].clone().requires_grad_(True)
    params = dict(fr.params, poses=poses)
    if correct_Rs is not None:
        correct_Rs = correct_Rs.clone().requires_grad_(True)
    xyz = g.xyz.clone().requires_grad_(True)
    scal = g.scales.clone().requires_grad_(True)
    rot = raw_rot.clone().requires_grad_(True)
    if lbs_delta is not None:
        lbs_delta.requires_grad_(True)
    m._xyz = xyz
    m._scaling = torch.log(scal)
    m._rotation = rot
    smpl_pts, world, bw, transforms, transl = m.coarse_deform_c2source(
        xyz[None], params, fr.big_params, t_vertices, lbs_weights=lbs_delta, correct_Rs=correct_Rs,
        return_transl=True)
    cov = m.get_covariance(mod, transforms.squeeze(0))
    # upstream gradients as the rasterizer would hand them back (dL/dmeans3D, dL/dcov3D), seeded
    g_world = torch.from_numpy(rng.standard_normal((P, 3))).float()
    g_cov = torch.from_numpy(rng.standard_normal((P, 6)) * 50.0).float()
    leaves = [xyz, scal, rot, poses] + ([lbs_delta] if lbs_delta is not None else []) + ([correct_Rs] if correct_Rs is not None else [])
    grads = list(torch.autograd.grad([world[0], cov], leaves, [g_world, g_cov]))
    d_poses = grads.pop(3)
    d_correct_Rs = grads.pop() if correct_Rs is not None else None
    with torch.no_grad():
        _, vert_ids = m.knn(t_vertices.float(), g.xyz[None].float())
    smpl_pts, world, bw, transforms, transl, cov = (t.detach() for t in (smpl_pts, world, bw, transforms, transl, cov))
    out = dict(P=P, seed=seed, mod=np.float32(mod), xyz=g.xyz.numpy(), scaling=g.scales.numpy(), rotation=raw_rot.numpy(),
               vert_ids=vert_ids.reshape(-1).numpy().astype(np.int32),
               smpl_pts=smpl_pts[0].numpy(), world=world[0].numpy(), bweights=bw[0].numpy(),
               transforms=transforms[0].numpy(), translation=transl[0].numpy(), cov3D=cov.numpy())
    out.update(g_world=g_world.numpy(), g_cov=g_cov.numpy(), d_xyz=grads[0].numpy(), d_scaling=grads[1].numpy(),
               d_rotation=grads[2].numpy(), d_poses=d_poses.numpy())
    if lbs_delta is not None:
        out["lbs_delta"] = lbs_delta[0].detach().numpy()
        out["d_lbs_delta"] = grads[3][0].numpy()
    if correct_Rs is not None:
        out["correct_Rs"] = correct_Rs[0].detach().numpy()
        out["d_correct_Rs"] = d_correct_Rs[0].numpy()
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), f"lbs_{name}.npz")
    np.savez_compressed(path, **out)
    print(name, "->", path, {k: getattr(v, "shape", v) for k, v in out.items()})


def make_frame_case(name, seed, with_correct_Rs, n_shapes=10):
    """Per-frame tables alone: the real reference's get_transform_params_torch (:998-1031,
    chain :965-995), batch_rodrigues (:1033-1064) and the blend-shape expressions of coarse_deform_c2source (:855-861,
    :870-871, :878-893) on one frame, plus autograd gradients w.r.t. poses / correct_Rs for seeded upstream gradients."""
    from moss_b200 import synth
    body = synth.body()
    GM, m = reference_model(body)
    fr = synth.frame_params(seed)
    rng = np.random.default_rng(1234 + seed)
    V = body.verts.shape[0]
    smpl = m.SMPL_NEUTRAL
    poses = fr.params["poses"].clone().requires_grad_(True)
    shapes = fr.params["shapes"][:, :n_shapes].clone()
    params = dict(fr.params, poses=poses, shapes=shapes)
    correct_Rs = None
    if with_correct_Rs:
        aa = torch.from_numpy(0.05 * rng.standard_normal((23, 3))).float()
        correct_Rs = GM.batch_rodrigues(aa).reshape(1, 23, 3, 3).clone().requires_grad_(True)
    # the statements of coarse_deform_c2source :878-893 (the function itself only returns per-Gaussian results)
    rot_mats = GM.batch_rodrigues(poses.view(-1, 3)).view([1, -1, 3, 3])
    if correct_Rs is not None:
        rot_mats_no_root = rot_mats[:, 1:]
        rot_mats_no_root = torch.matmul(rot_mats_no_root.reshape(-1, 3, 3), correct_Rs.reshape(-1, 3, 3)).reshape(-1, 23, 3, 3)
        rot_mats = torch.cat([rot_mats[:, 0:1], rot_mats_no_root], dim=1)
    ident = torch.eye(3)
    pose_feature = (rot_mats[:, 1:, :, :] - ident).view([1, -1])
    pose_offsets = torch.matmul(pose_feature.unsqueeze(1), smpl["posedirs"].view(V * 3, -1).transpose(1, 0).unsqueeze(0)).view(1, -1, 3)
    shapedirs = smpl["shapedirs"][..., :shapes.shape[-1]]
    shape_offset = torch.matmul(shapedirs.unsqueeze(0), torch.reshape(shapes, (1, 1, -1, 1))).squeeze(-1)
    A, R, Th, joints = GM.get_transform_params_torch(smpl, params, rot_mats=rot_mats)
    g_A = torch.from_numpy(rng.standard_normal((1, 24, 4, 4))).float()
    g_A[:, :, 3, :] = 0          # the bottom row of A is constant
    g_po = torch.from_numpy(rng.standard_normal((1, V, 3))).float()
    leaves = [poses] + ([correct_Rs] if correct_Rs is not None else [])
    grads = torch.autograd.grad([A, pose_offsets], leaves, [g_A, g_po])
    out = dict(seed=seed, n_shapes=n_shapes, poses=poses.detach().numpy(), shapes=shapes.numpy(),
               A=A[0].detach().numpy(), joints=joints[0].detach().numpy(), rot_mats=rot_mats[0].detach().numpy(),
               pose_offsets=pose_offsets[0].detach().numpy(), shape_offset=shape_offset[0].detach().numpy(),
               g_A=g_A[0].numpy(), g_pose_offsets=g_po[0].numpy(), d_poses=grads[0].numpy())
    if correct_Rs is not None:
        out["correct_Rs"] = correct_Rs[0].detach().numpy()
        out["d_correct_Rs"] = grads[1][0].numpy()
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), f"smpl_frame_{name}.npz")
    np.savez_compressed(path, **out)
    print(name, "->", path, {k: getattr(v, "shape", v) for k, v in out.items()})


if __name__ == "__main__":
    make_frame_case("plain", 0, False)
    make_frame_case("correct", 2, True)
    make_frame_case("shapes6", 3, True, n_shapes=6)
    make_case("plain", 1500, 0, False, False, 1.0)
    make_case("delta", 1200, 1, True, False, 1.0)
    make_case("correct", 1000, 2, True, True, 0.8)
