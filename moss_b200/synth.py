"""Deterministic synthetic scenes shaped like MOSS's workloads (no SMPL pkl / datasets offline).

  * body():  a 24-joint humanoid with SMPL's kinematic tree (nets/mlp_delta_body_pose.py:41) and 6 890
    surface points on capsule limbs with <=4 non-zero skinning weights per point ("SMPL topology").
  * gaussians(): P Gaussians = body points + jittered clones (mimics densify split/clone), scales from the
    mean 3-NN distance (scene/gaussian_model.py:185-186), random rotations, opacities, SH (degree-3 storage).
  * camera()/ring_cameras(): ZJU-like cameras built exactly like scene/cameras.py:60-65 with
    utils/graphics_utils.py:39-50 (getWorld2View2) and :83-103 (getProjectionMatrix_refine).
  * stress_scene(): 1 M-Gaussian generic 3DGS scene at 1920x1080 (BASELINE.json configs[4]).

Everything is generated on the CPU with a seeded numpy generator (seed 3407, utils/general_utils.py:141)
and returned as CPU float32 tensors; callers move them to the GPU.
"""
import math
from typing import NamedTuple

import numpy as np
import torch

SEED = 3407
SMPL_PARENTS = [-1, 0, 0, 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 9, 9, 12, 13, 14, 16, 17, 18, 19, 20, 21]
N_SMPL_VERTS = 6890

# rest-pose joint positions of a ~1.7 m humanoid (metres; y up), SMPL joint order
_REST_JOINTS = np.array([
    [0.00, 0.95, 0.00],   # 0 pelvis
    [0.09, 0.87, 0.00],   # 1 l_hip
    [-0.09, 0.87, 0.00],  # 2 r_hip
    [0.00, 1.06, 0.00],   # 3 spine1
    [0.10, 0.49, 0.00],   # 4 l_knee
    [-0.10, 0.49, 0.00],  # 5 r_knee
    [0.00, 1.19, 0.00],   # 6 spine2
    [0.10, 0.09, 0.00],   # 7 l_ankle
    [-0.10, 0.09, 0.00],  # 8 r_ankle
    [0.00, 1.25, 0.00],   # 9 spine3
    [0.11, 0.03, 0.12],   # 10 l_foot
    [-0.11, 0.03, 0.12],  # 11 r_foot
    [0.00, 1.46, 0.00],   # 12 neck
    [0.08, 1.37, 0.00],   # 13 l_collar
    [-0.08, 1.37, 0.00],  # 14 r_collar
    [0.00, 1.58, 0.00],   # 15 head
    [0.18, 1.40, 0.00],   # 16 l_shoulder
    [-0.18, 1.40, 0.00],  # 17 r_shoulder
    [0.44, 1.40, 0.00],   # 18 l_elbow
    [-0.44, 1.40, 0.00],  # 19 r_elbow
    [0.69, 1.40, 0.00],   # 20 l_wrist
    [-0.69, 1.40, 0.00],  # 21 r_wrist
    [0.78, 1.40, 0.00],   # 22 l_hand
    [-0.78, 1.40, 0.00],  # 23 r_hand
], dtype=np.float64)
_BONE_RADIUS = {0: 0.13, 3: 0.14, 6: 0.15, 9: 0.14, 12: 0.06, 15: 0.10, 1: 0.085, 2: 0.085, 4: 0.06, 5: 0.06,
                7: 0.045, 8: 0.045, 10: 0.04, 11: 0.04, 13: 0.07, 14: 0.07, 16: 0.055, 17: 0.055, 18: 0.045,
                19: 0.045, 20: 0.035, 21: 0.035, 22: 0.03, 23: 0.03}


class Body(NamedTuple):
    verts: torch.Tensor      # [6890,3] rest-pose surface points
    joints: torch.Tensor     # [24,3]
    parents: torch.Tensor    # [24] int64
    weights: torch.Tensor    # [6890,24] rows sum to 1, <=4 non-zeros
    shapedirs: torch.Tensor  # [6890,3,10]
    posedirs: torch.Tensor   # [207, 6890*3]
    J_regressor: torch.Tensor  # [24,6890]


def body(seed=SEED) -> Body:
    rng = np.random.default_rng(seed)
    J = _REST_JOINTS
    segs = [(p, j) for j, p in enumerate(SMPL_PARENTS) if p >= 0]
    lens = np.array([np.linalg.norm(J[j] - J[p]) * (_BONE_RADIUS[j] + 0.02) for p, j in segs])
    counts = np.floor(lens / lens.sum() * N_SMPL_VERTS).astype(int)
    counts[0] += N_SMPL_VERTS - counts.sum()
    pts = []
    for (p, j), n in zip(segs, counts):
        a, b = J[p], J[j]
        axis = (b - a) / (np.linalg.norm(b - a) + 1e-12)
        tmp = np.array([1.0, 0, 0]) if abs(axis[0]) < 0.9 else np.array([0, 1.0, 0])
        u = np.cross(axis, tmp); u /= np.linalg.norm(u)
        w = np.cross(axis, u)
        t = rng.uniform(-0.08, 1.08, n)
        ang = rng.uniform(0, 2 * math.pi, n)
        rad = _BONE_RADIUS[j] * (0.92 + 0.08 * rng.standard_normal(n))
        pts.append(a[None] + t[:, None] * (b - a)[None] + rad[:, None] * (np.cos(ang)[:, None] * u[None] + np.sin(ang)[:, None] * w[None]))
    verts = np.concatenate(pts, 0)
    # skinning weights: softmax over negative bone distance, top-4, renormalised
    d = np.stack([_seg_dist(verts, J[SMPL_PARENTS[j]] if SMPL_PARENTS[j] >= 0 else J[j], J[j]) for j in range(24)], 1)
    logit = -d / 0.04
    logit -= logit.max(1, keepdims=True)
    wgt = np.exp(logit)
    kth = np.sort(wgt, 1)[:, -4][:, None]
    wgt = np.where(wgt >= kth, wgt, 0.0)
    wgt /= wgt.sum(1, keepdims=True)
    shapedirs = 0.01 * rng.standard_normal((N_SMPL_VERTS, 3, 10))
    posedirs = 0.002 * rng.standard_normal((207, N_SMPL_VERTS * 3))
    # joint regressor: each joint = mean of its 32 nearest surface points shifted back onto the joint
    Jreg = np.zeros((24, N_SMPL_VERTS))
    for j in range(24):
        nn = np.argsort(np.linalg.norm(verts - J[j][None], axis=1))[:32]
        Jreg[j, nn] = 1.0 / 32
    f = lambda a, dt=torch.float32: torch.from_numpy(np.ascontiguousarray(a)).to(dt)  # noqa: E731
    return Body(f(verts), f(Jreg @ verts), torch.tensor(SMPL_PARENTS, dtype=torch.int64), f(wgt), f(shapedirs),
                f(posedirs), f(Jreg))


def _seg_dist(p, a, b):
    ab = b - a
    den = float(ab @ ab)
    if den < 1e-12:
        return np.linalg.norm(p - a[None], axis=1)
    t = np.clip(((p - a[None]) @ ab) / den, 0, 1)
    return np.linalg.norm(p - (a[None] + t[:, None] * ab[None]), axis=1)


class Gaussians(NamedTuple):
    xyz: torch.Tensor        # [P,3]
    scales: torch.Tensor     # [P,3] (activated, i.e. exp of the stored log-scale)
    rotations: torch.Tensor  # [P,4] normalised (r,x,y,z)
    opacities: torch.Tensor  # [P,1] in (0.05, 0.99)
    shs: torch.Tensor        # [P,16,3]
    vert_ids: torch.Tensor   # [P] nearest body vertex (the clone's parent)


def gaussians(P, b: Body = None, seed=SEED) -> Gaussians:
    rng = np.random.default_rng(seed + 1)
    b = b or body(seed)
    V = b.verts.numpy().astype(np.float64)
    nv = V.shape[0]
    ids = np.concatenate([np.arange(min(P, nv)), rng.integers(0, nv, max(P - nv, 0))])[:P]
    xyz = V[ids].copy()
    if P > nv:
        xyz[nv:] += 0.012 * rng.standard_normal((P - nv, 3))
    # scale init: sqrt(mean squared distance to the 3 nearest neighbours) (simple-knn distCUDA2 rule)
    dist2 = _mean_knn_dist2(xyz, 3)
    base = np.sqrt(np.maximum(dist2, 1e-7))
    scales = base[:, None] * rng.uniform(0.5, 1.5, (P, 3))
    q = rng.standard_normal((P, 4)); q /= np.linalg.norm(q, axis=1, keepdims=True)
    logit = rng.uniform(-2.9, 4.5, (P, 1))
    opac = np.clip(1 / (1 + np.exp(-logit)), 0.05, 0.99)
    sh = np.zeros((P, 16, 3))
    sh[:, 0] = (rng.uniform(0, 1, (P, 3)) - 0.5) / 0.28209479177387814
    sh[:, 1:] = 0.05 * rng.standard_normal((P, 15, 3))
    f = lambda a: torch.from_numpy(np.ascontiguousarray(a)).float()  # noqa: E731
    return Gaussians(f(xyz), f(scales), f(q), f(opac), f(sh), torch.from_numpy(ids).long())


def _mean_knn_dist2(x, k):
    from scipy.spatial import cKDTree
    d, _ = cKDTree(x).query(x, k=k + 1)
    return (d[:, 1:] ** 2).mean(1)


class Camera(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    world_view_transform: torch.Tensor   # [4,4] = W2C^T
    full_proj_transform: torch.Tensor    # [4,4]
    camera_center: torch.Tensor          # [3]
    K: torch.Tensor
    R: torch.Tensor                      # as stored by the reference's Camera (C2W rotation)
    T: torch.Tensor


def camera(eye, target, W, H, fx, fy=None, cx=None, cy=None, up=(0.0, 1.0, 0.0), znear=0.001, zfar=1000.0) -> Camera:
    """Pinhole camera looking from `eye` at `target` (OpenCV axes: x right, y down, z forward)."""
    fy = fx if fy is None else fy
    cx = W / 2 if cx is None else cx
    cy = H / 2 if cy is None else cy
    eye, target, up = (np.asarray(a, dtype=np.float64) for a in (eye, target, up))
    zc = target - eye; zc /= np.linalg.norm(zc)
    xc = np.cross(zc, up); xc /= np.linalg.norm(xc)     # right
    yc = np.cross(zc, xc)                               # down
    Rw2c = np.stack([xc, yc, zc], 0)
    t = -Rw2c @ eye
    R = Rw2c.T                                           # reference stores R = W2C rotation transposed
    # getWorld2View2 (utils/graphics_utils.py:39-50) with translate 0, scale 1
    Rt = np.zeros((4, 4)); Rt[:3, :3] = R.T; Rt[:3, 3] = t; Rt[3, 3] = 1.0
    Rt = np.float32(np.linalg.inv(np.linalg.inv(Rt)))
    wvt = torch.tensor(Rt).transpose(0, 1).contiguous()
    K = torch.tensor([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], dtype=torch.float32)
    Pm = torch.zeros(4, 4)                               # getProjectionMatrix_refine (:83-103)
    Pm[0, 0] = 2 * K[0, 0] / W; Pm[0, 1] = 2 * K[0, 1] / W; Pm[0, 2] = -1 + 2 * (K[0, 2] / W)
    Pm[1, 1] = 2 * K[1, 1] / H; Pm[1, 2] = -1 + 2 * (K[1, 2] / H)
    Pm[2, 2] = (zfar + znear) / (zfar - znear); Pm[2, 3] = -2 * zfar * znear / (zfar - znear); Pm[3, 2] = 1.0
    proj = Pm.transpose(0, 1)
    full = (wvt.unsqueeze(0).bmm(proj.unsqueeze(0))).squeeze(0).contiguous()
    center = wvt.inverse()[3, :3].contiguous()
    return Camera(H, W, W / (2 * fx), H / (2 * fy), wvt, full, center, K,
                  torch.from_numpy(R).float(), torch.from_numpy(t).float())


def ring_cameras(n=23, W=512, H=512, radius=3.0, height=1.0, target=(0.0, 0.95, 0.0), seed=SEED):
    """ZJU-MoCap-like ring of `n` cameras; fx=fy~=540 px at 512^2 (scaled with W), principal point jittered."""
    rng = np.random.default_rng(seed + 2)
    cams = []
    fx = 540.0 * W / 512.0
    for i in range(n):
        a = 2 * math.pi * i / n
        eye = (radius * math.sin(a), height, radius * math.cos(a))
        cams.append(camera(eye, target, W, H, fx, fx, W / 2 + rng.uniform(-10, 10), H / 2 + rng.uniform(-10, 10),
                           up=(0.0, -1.0, 0.0)))
    return cams


class StressScene(NamedTuple):
    g: Gaussians
    cam: Camera


def stress_scene(P=1_000_000, W=1920, H=1080, seed=SEED) -> StressScene:
    """Generic 3DGS-like scene: 64 blobs in a 10 m box, 2-20 m in front of a 60 deg camera, log-normal scales."""
    rng = np.random.default_rng(seed + 3)
    centers = np.stack([rng.uniform(-5, 5, 64), rng.uniform(-2.5, 2.5, 64), rng.uniform(2, 20, 64)], 1)
    sig = rng.uniform(0.3, 1.2, 64)
    which = rng.integers(0, 64, P)
    xyz = centers[which] + sig[which][:, None] * rng.standard_normal((P, 3))
    scales = np.exp(math.log(0.02) + 0.5 * rng.standard_normal((P, 3)))
    q = rng.standard_normal((P, 4)); q /= np.linalg.norm(q, axis=1, keepdims=True)
    opac = np.clip(1 / (1 + np.exp(-rng.uniform(-2.9, 4.5, (P, 1)))), 0.05, 0.99)
    sh = np.zeros((P, 16, 3), dtype=np.float32)
    sh[:, 0] = (rng.uniform(0, 1, (P, 3)) - 0.5) / 0.28209479177387814
    sh[:, 1:] = 0.05 * rng.standard_normal((P, 15, 3)).astype(np.float32)
    f = lambda a: torch.from_numpy(np.ascontiguousarray(a)).float()  # noqa: E731
    fx = W / (2 * math.tan(math.radians(60) / 2))
    cam = camera((0.0, 0.0, 0.0), (0.0, 0.0, 1.0), W, H, fx, fx, up=(0.0, -1.0, 0.0))
    return StressScene(Gaussians(f(xyz), f(scales), f(q), f(opac), f(sh), torch.zeros(P, dtype=torch.long)), cam)


def settings_kwargs(cam: Camera, bg, sh_degree=3, scale_modifier=1.0, device=None):
    """Keyword arguments for GaussianRasterizationSettings, as gaussian_renderer/__init__.py:36-52 builds them."""
    to = (lambda t: t.to(device)) if device is not None else (lambda t: t)
    return dict(image_height=int(cam.image_height), image_width=int(cam.image_width), tanfovx=cam.tanfovx,
                tanfovy=cam.tanfovy, bg=to(bg), scale_modifier=scale_modifier, viewmatrix=to(cam.world_view_transform),
                projmatrix=to(cam.full_proj_transform), sh_degree=sh_degree, campos=to(cam.camera_center),
                prefiltered=False, debug=False)


class Frame(NamedTuple):
    params: dict       # target pose: poses [1,72], shapes [1,10], R [3,3], Th [1,3]  (scene/dataset_readers.py:420-426)
    big_params: dict   # canonical "big pose" (scene/dataset_readers.py:583-590)


def frame_params(seed=0) -> Frame:
    """Random SMPL frame parameters shaped like the reference's `smpl_param` / `big_pose_smpl_param` dicts."""
    rng = np.random.default_rng(SEED + 17 + seed)
    poses = 0.35 * rng.standard_normal((1, 72))
    poses[0, :3] = 0.2 * rng.standard_normal(3)
    rh = 0.4 * rng.standard_normal(3)
    ang = np.linalg.norm(rh)
    k = rh / ang
    K = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    Rg = np.eye(3) + math.sin(ang) * K + (1 - math.cos(ang)) * (K @ K)      # cv2.Rodrigues(Rh)
    f = lambda a: torch.from_numpy(np.ascontiguousarray(a)).float()  # noqa: E731
    params = dict(poses=f(poses), shapes=f(0.5 * rng.standard_normal((1, 10))), R=f(Rg),
                  Th=f(0.3 * rng.standard_normal((1, 3))))
    big = np.zeros((1, 72))
    big[0, 5] = 45 / 180 * math.pi; big[0, 8] = -45 / 180 * math.pi
    big[0, 23] = -30 / 180 * math.pi; big[0, 26] = 30 / 180 * math.pi
    big_params = dict(poses=f(big), shapes=torch.zeros(1, 10), R=torch.eye(3), Th=torch.zeros(1, 3))
    return Frame(params, big_params)
