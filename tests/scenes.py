"""Seeded test scenes (CPU tensors) shared by the CPU and GPU suites."""
import math

import numpy as np
import torch

from moss_b200 import synth


def _cov_from_scale_rot_python(scales, rotations, mod=1.0, transforms=None):
    """MOSS's python covariance (scene/gaussian_model.py:37-44, utils/general_utils.py:65-118):
    normalised quaternion, L = R diag(s*mod), Sigma = L L^T, optional Sigma' = T Sigma T^T, 6-float upper triangle."""
    q = rotations / rotations.norm(dim=1, keepdim=True)
    r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = torch.stack([
        1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
        2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
        2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], 1).view(-1, 3, 3)
    L = R * (scales * mod)[:, None, :]
    S = L @ L.transpose(1, 2)
    if transforms is not None:
        S = transforms @ S @ transforms.transpose(1, 2)
    return torch.stack([S[:, 0, 0], S[:, 0, 1], S[:, 0, 2], S[:, 1, 1], S[:, 1, 2], S[:, 2, 2]], 1).contiguous()


def make_scene(name):
    """-> dict(means3D, shs, colors_precomp, opacities, scales, rotations, cov3D_precomp, cam, bg, sh_degree,
    scale_modifier).  Exactly one of shs/colors_precomp and one of (scales,rotations)/cov3D_precomp is not None."""
    return make_scene_from_spec(SCENES[name])


def make_scene_from_spec(spec):
    """The scene an entry of SCENES describes (also used by the randomised specs of tests/test_zy_fuzz_gpu.py)."""
    P, W, H = spec["P"], spec["W"], spec["H"]
    if spec.get("stress"):
        sc = synth.stress_scene(P, W, H)
        g, cam = sc.g, sc.cam
    else:
        g = synth.gaussians(P, seed=synth.SEED + spec.get("seed", 0))
        cam = synth.ring_cameras(23, W, H, radius=spec.get("radius", 3.0))[spec.get("cam", 0)]
    d = dict(means3D=g.xyz.clone(), shs=g.shs.clone(), colors_precomp=None, opacities=g.opacities.clone(),
             scales=g.scales.clone(), rotations=g.rotations.clone(), cov3D_precomp=None, cam=cam,
             bg=torch.tensor(spec.get("bg", [0.0, 0.0, 0.0])), sh_degree=spec.get("D", 3),
             scale_modifier=spec.get("mod", 1.0))
    rng = np.random.default_rng(1234 + P)
    if spec.get("raw_quat"):        # un-normalised quaternions exercise DGR forward.cu:127
        d["rotations"] = d["rotations"] * torch.from_numpy(rng.uniform(0.6, 1.5, (P, 1))).float()
    if spec.get("scale_mul"):
        d["scales"] = d["scales"] * spec["scale_mul"]
    if spec.get("precomp_color"):
        d["colors_precomp"] = torch.from_numpy(rng.uniform(0, 1, (P, 3))).float()
        d["shs"] = None
    if spec.get("precomp_cov"):     # MOSS's default path (compute_cov3D_python is always on)
        T = None
        if spec.get("transforms"):
            A = torch.from_numpy(rng.standard_normal((P, 3, 3)) * 0.05).float()
            T = torch.eye(3)[None] + A
        d["cov3D_precomp"] = _cov_from_scale_rot_python(d["scales"], d["rotations"], d["scale_modifier"], T)
        d["scales"] = None
        d["rotations"] = None
    if spec.get("behind"):          # push some Gaussians behind / onto the near plane (behind = 1.0: all of them, R = 0)
        m = d["means3D"]
        frac = 0.1 if spec["behind"] is True else float(spec["behind"])
        idx = torch.arange(P) if frac >= 1.0 else torch.from_numpy(rng.integers(0, P, max(int(P * frac), 1)))
        m[idx] = m[idx] + cam.camera_center.clone()[None] * 1.2
    if spec.get("sh_M") and d["shs"] is not None:      # stored SH coefficients per Gaussian (1, 4, 9 or 16)
        d["shs"] = d["shs"][:, :spec["sh_M"]].contiguous()
    if spec.get("opacity_edges"):   # exactly transparent and exactly opaque Gaussians
        idx = torch.from_numpy(rng.integers(0, P, max(P // 8, 1)))
        d["opacities"][idx[: len(idx) // 2]] = 0.0
        d["opacities"][idx[len(idx) // 2:]] = 1.0
    return d


SCENES = {
    # name: P, W, H, options
    "tiny": dict(P=300, W=64, H=48, radius=2.2),
    "small": dict(P=2000, W=128, H=128),
    "small_cov": dict(P=2000, W=128, H=128, precomp_cov=True, transforms=True),
    "small_color": dict(P=2000, W=128, H=128, precomp_color=True, bg=[1.0, 1.0, 1.0]),
    "small_d0": dict(P=1500, W=128, H=96, D=0),
    "small_d1": dict(P=1500, W=128, H=96, D=1),
    "small_d2": dict(P=1500, W=128, H=96, D=2, mod=0.7),
    "ragged": dict(P=3000, W=200, H=120, raw_quat=True, bg=[0.2, 0.5, 0.9], behind=True, cam=5),
    "big_splats": dict(P=1200, W=160, H=160, scale_mul=6.0, cam=3),
    "zju": dict(P=20000, W=512, H=512, precomp_cov=True, transforms=True),
    "zju_sr": dict(P=20000, W=512, H=512),
    "monocap": dict(P=100000, W=1024, H=1024, precomp_cov=True, transforms=True),
    # edge-on view: > 11 000 instances in one tile (beyond the shared-memory sort; exercises tile_sort_large)
    "monocap_edge": dict(P=100000, W=1024, H=1024, precomp_cov=True, transforms=True, cam=6),
    "stress": dict(P=1000000, W=1920, H=1080, stress=True),
    "stress_small": dict(P=60000, W=480, H=270, stress=True),
}
CPU_SCENES = ["tiny", "small", "small_cov", "small_color", "small_d0", "small_d1", "small_d2", "ragged", "big_splats"]


def upstream_grads(W, H, seed=7, depth=True):
    g = torch.Generator().manual_seed(seed)
    N = W * H
    gc = torch.randn((3, H, W), generator=g) / N
    ga = torch.randn((1, H, W), generator=g) / N
    gd = torch.randn((1, H, W), generator=g) / N if depth else torch.zeros((1, H, W))
    return gc, gd, ga


def oracle_settings(d):
    from oracle import raster_oracle as O
    return O.Settings(**synth.settings_kwargs(d["cam"], d["bg"], d["sh_degree"], d["scale_modifier"]))
