"""CPU: the LBS / covariance oracle restatement reproduces what the REAL reference functions computed
(tests/golden/lbs_*.npz, made by tests/golden/make_lbs_golden.py in the build container)."""
import os

import numpy as np
import pytest
import torch

from oracle import lbs_oracle as L
from tests.lbs_cases import GOLD, LBS_CASES, load_case

TOL = 2e-6      # fp32; both sides are the same torch ops up to matmul association


def _close(a, b, tol=TOL):
    scale = max(float(b.abs().max()), 1.0)
    assert float((a - b).abs().max()) <= tol * scale, float((a - b).abs().max()) / scale


@pytest.mark.parametrize("name", LBS_CASES)
def test_forward_matches_reference(name):
    d, ft, body = load_case(name)
    vid = L.nearest_vertex(body.verts, d["xyz"])
    assert torch.equal(vid.int(), d["vert_ids"])
    sk, cov = L.skin_and_covariance(d["xyz"], vid, body.weights, ft, d["scaling"], d["rotation"], float(d["mod"]),
                                    d.get("lbs_delta"), return_transl=True)
    _close(sk.bweights, d["bweights"])
    _close(sk.smpl_pts, d["smpl_pts"])
    _close(sk.world_pts, d["world"])
    _close(sk.transforms, d["transforms"])
    _close(sk.translation, d["translation"])
    _close(cov, d["cov3D"], 1e-5)


@pytest.mark.parametrize("name", LBS_CASES)
def test_backward_matches_reference(name):
    d, ft, body = load_case(name)
    xyz = d["xyz"].clone().requires_grad_(True)
    sc = d["scaling"].clone().requires_grad_(True)
    rot = d["rotation"].clone().requires_grad_(True)
    leaves = [xyz, sc, rot]
    delta = None
    if "lbs_delta" in d:
        delta = d["lbs_delta"].clone().requires_grad_(True)
        leaves.append(delta)
    sk, cov = L.skin_and_covariance(xyz, d["vert_ids"].long(), body.weights, ft, sc, rot, float(d["mod"]), delta)
    g = torch.autograd.grad([sk.world_pts, cov], leaves, [d["g_world"], d["g_cov"]])
    for got, key in zip(g, ["d_xyz", "d_scaling", "d_rotation", "d_lbs_delta"]):
        rel = float((got - d[key]).norm() / d[key].norm())
        assert rel < 1e-4, (key, rel)


FRAME_CASES = ["plain", "correct", "shapes6"]


def load_frame_case(name):
    z = np.load(os.path.join(GOLD, f"smpl_frame_{name}.npz"))
    return {k: torch.from_numpy(np.asarray(z[k])) if getattr(z[k], "ndim", 0) > 0 else z[k].item() for k in z.files}


@pytest.mark.parametrize("name", FRAME_CASES)
def test_oracle_frame_tables_match_reference_golden(name):
    """oracle/lbs_oracle.py's per-frame tables (and their autograd gradients w.r.t. poses / correct_Rs) against the real
    reference run by tests/golden/make_lbs_golden.py::make_frame_case."""
    from oracle import lbs_oracle as L
    from moss_b200 import synth
    from tests.lbs_cases import oracle_smpl
    d = load_frame_case(name)
    body = synth.body()
    fr = synth.frame_params(int(d["seed"]))
    poses = d["poses"].clone().requires_grad_(True)
    cr = d["correct_Rs"].reshape(1, 23, 3, 3).clone().requires_grad_(True) if "correct_Rs" in d else None
    params = dict(fr.params, poses=poses, shapes=d["shapes"])
    ft = L.frame_tables(oracle_smpl(body), params, fr.big_params, cr)
    for got, key in ((ft.A_tgt, "A"), (ft.poseoff_tgt, "pose_offsets"), (ft.shapeoff, "shape_offset")):
        assert float((got.detach() - d[key]).abs().max()) <= 2e-6 * max(1.0, float(d[key].abs().max())), key
    leaves = [poses] + ([cr] if cr is not None else [])
    grads = torch.autograd.grad([ft.A_tgt, ft.poseoff_tgt], leaves, [d["g_A"], d["g_pose_offsets"]])
    assert float((grads[0] - d["d_poses"]).norm() / d["d_poses"].norm()) < 1e-5
    if cr is not None:
        assert float((grads[1][0] - d["d_correct_Rs"]).norm() / d["d_correct_Rs"].norm()) < 1e-5
