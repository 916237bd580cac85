"""GPU: the per-frame SMPL table kernels (moss_smpl_frame_forward / _backward through the C ABI)
against
  * tests/golden/smpl_frame_*.npz -- A, joints, rotations, pose / shape offsets AND autograd gradients w.r.t. poses and
    correct_Rs of the REAL reference functions (get_transform_params_torch, batch_rodrigues, the blend-shape
    expressions of coarse_deform_c2source; tests/golden/make_lbs_golden.py::make_frame_case),
  * tests/golden/lbs_*.npz::d_poses / d_correct_Rs -- the real reference's coarse_deform_c2source + get_covariance
    differentiated down to the pose parameters: frame kernel -> skin kernel -> both backwards, end to end
    (pins d_A_tgt / d_poseoff_tgt of the skin backward on the reference, not on the oracle's restatement).
Tolerances (fp32, different summation order than torch's matmuls): forward 1e-5 of the tensor's max-abs, gradients
1e-4 norm-relative (north_star allows 1e-3)."""
import ctypes as C

import pytest
import torch

from moss_b200 import _abi, skinning as K, synth
from tests.lbs_cases import LBS_CASES, load_case
from tests.test_lbs_oracle_golden import FRAME_CASES, load_frame_case

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _smpl(body, dev=DEV):
    V = body.verts.shape[0]
    return K.Smpl(body.verts.to(dev), body.shapedirs.to(dev), body.posedirs.t().reshape(V, 3, 207).contiguous().to(dev),
                  body.J_regressor.to(dev), body.weights.to(dev), body.parents.to(dev))


def _close(a, b, tol=1e-5):
    a, b = a.detach().cpu().double(), b.detach().cpu().double().reshape(a.shape)
    err = float((a - b).abs().max()) / max(float(b.abs().max()), 1.0)
    assert err <= tol, err


def _rel(a, b):
    a, b = a.detach().cpu().double(), b.detach().cpu().double().reshape(a.shape)
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


@pytest.mark.parametrize("name", FRAME_CASES)
def test_frame_forward_and_backward_match_reference_golden(name):
    d = load_frame_case(name)
    smpl = _smpl(synth.body())
    poses = d["poses"].to(DEV).requires_grad_(True)
    cr = d["correct_Rs"].to(DEV).requires_grad_(True) if "correct_Rs" in d else None
    A, po, so = K.smpl_frame(smpl, poses, d["shapes"].to(DEV), cr)
    _close(A, d["A"]); _close(po, d["pose_offsets"]); _close(so, d["shape_offset"])
    assert A.shape == (24, 4, 4) and torch.equal(A[:, 3].cpu(), torch.tensor([0.0, 0, 0, 1]).expand(24, 4))
    torch.autograd.backward([A, po], [d["g_A"].to(DEV), d["g_pose_offsets"].to(DEV)])
    assert poses.grad.shape == poses.shape and _rel(poses.grad, d["d_poses"]) < 1e-4
    if cr is not None:
        assert cr.grad.shape == cr.shape and _rel(cr.grad, d["d_correct_Rs"]) < 1e-4


def test_raw_abi_optional_outputs_joints_rotations_and_reproducibility():
    """rot_mats / joints through the raw entry point; NULL poseoff skips the posedirs pass; two calls give the same bits
    (block-ordered partial sums, no float atomics); the scratch is left clean for the next call."""
    d = load_frame_case("correct")
    smpl = _smpl(synth.body())
    lib = _abi.load()
    V = smpl.v_template.shape[0]
    poses, shapes, cr = d["poses"].reshape(-1).to(DEV), d["shapes"].reshape(-1).to(DEV), d["correct_Rs"].reshape(-1).to(DEV)
    scratch = torch.zeros(lib.moss_smpl_frame_scratch_bytes(V), dtype=torch.uint8, device=DEV)
    parents = smpl.parents.to(torch.int32)                 # kept referenced: the struct only holds its address
    f = K._fill_frame(smpl, poses, shapes, cr, scratch, parents)
    outs = []
    for rep in range(2):
        A, R, J = torch.empty(24, 16, device=DEV), torch.empty(24, 9, device=DEV), torch.empty(24, 3, device=DEV)
        po, so = torch.full((V, 3), 7.0, device=DEV), torch.empty(V, 3, device=DEV)
        st = lib.moss_smpl_frame_forward(None, C.byref(f), A.data_ptr(), R.data_ptr(), J.data_ptr(),
                                         po.data_ptr() if rep == 0 else None, so.data_ptr())
        assert st == _abi.MOSS_OK
        torch.cuda.synchronize()
        outs.append((A, R, J, so))
        if rep == 0:
            _close(po, d["pose_offsets"])
        else:
            assert float(po.min()) == 7.0          # untouched
    for x, y in zip(*outs):
        assert torch.equal(x, y)
    _close(outs[0][1], d["rot_mats"]); _close(outs[0][2], d["joints"])
    assert int(scratch[:16].view(torch.int32).abs().sum()) == 0
    # backward twice: identical bits, d_correct_Rs optional
    A, J = outs[0][0], outs[0][2]
    gA, gpo = d["g_A"].reshape(-1).to(DEV), d["g_pose_offsets"].reshape(-1).to(DEV)
    res = []
    for rep in range(2):
        dp, dc = torch.empty(72, device=DEV), torch.empty(23 * 9, device=DEV)
        st = lib.moss_smpl_frame_backward(None, C.byref(f), A.data_ptr(), J.data_ptr(), gA.data_ptr(), gpo.data_ptr(),
                                          dp.data_ptr(), dc.data_ptr() if rep == 0 else None)
        assert st == _abi.MOSS_OK
        torch.cuda.synchronize()
        res.append(dp)
    assert torch.equal(res[0], res[1]) and _rel(res[0], d["d_poses"]) < 1e-4
    # gradient through A alone / through the pose offsets alone add up to the joint gradient
    parts = []
    for ga, gp in ((gA, None), (None, gpo)):
        dp = torch.empty(72, device=DEV)
        assert lib.moss_smpl_frame_backward(None, C.byref(f), A.data_ptr(), J.data_ptr(), ga.data_ptr() if ga is not None else None,
                                            gp.data_ptr() if gp is not None else None, dp.data_ptr(), None) == _abi.MOSS_OK
        parts.append(dp)
    torch.cuda.synchronize()
    assert _rel(parts[0] + parts[1], res[0]) < 1e-5
    assert int(scratch[:16].view(torch.int32).abs().sum()) == 0      # every ticket back to zero


def test_frame_tables_match_oracle_and_cache():
    from oracle import lbs_oracle as L
    from tests.lbs_cases import oracle_smpl
    body = synth.body()
    smpl = _smpl(body)
    fr = synth.frame_params(3)
    cr = L.batch_rodrigues(0.05 * torch.randn(23, 3, generator=torch.Generator().manual_seed(1))).reshape(1, 23, 3, 3)
    todev = lambda p: {k: v.to(DEV) for k, v in p.items()}  # noqa: E731
    params, big = todev(fr.params), todev(fr.big_params)
    cache = K.FrameCache()
    a = K.frame_tables(smpl, params, big, cr.to(DEV), cache=cache)
    a2 = K.frame_tables(smpl, params, big, cr.to(DEV), cache=cache)
    a3 = K.frame_tables(smpl, params, big, cr.to(DEV))
    b = L.frame_tables(oracle_smpl(body), fr.params, fr.big_params, cr)
    for x, x2, x3, y in zip(a, a2, a3, b):
        _close(x, y, 2e-6)
        assert torch.equal(x, x2) and torch.equal(x, x3)
    assert a.A_big.data_ptr() == a2.A_big.data_ptr()          # the big-pose half came from the cache
    big["poses"].add_(0.01)                                    # in-place edit invalidates the cache
    a4 = K.frame_tables(smpl, params, big, cr.to(DEV), cache=cache)
    assert not torch.equal(a4.A_big, a.A_big)


@pytest.mark.parametrize("name", LBS_CASES)
def test_pose_gradients_end_to_end_match_reference_golden(name):
    """poses / correct_Rs -> frame kernel -> skin kernel -> (means3D, cov3D); the backward of both kernels against the
    real reference's autograd through coarse_deform_c2source + get_covariance."""
    d, _, body = load_case(name)
    smpl = _smpl(body)
    fr = synth.frame_params(int(d["seed"]))
    todev = lambda p: {k: v.to(DEV) for k, v in p.items()}  # noqa: E731
    params, big = todev(fr.params), todev(fr.big_params)
    params["poses"].requires_grad_(True)
    cr = d["correct_Rs"].reshape(1, 23, 3, 3).to(DEV).requires_grad_(True) if "correct_Rs" in d else None
    ft = K.frame_tables(smpl, params, big, cr)
    leaf = lambda t: t.to(DEV).requires_grad_(True)  # noqa: E731
    xyz, sc, rot = leaf(d["xyz"]), leaf(d["scaling"]), leaf(d["rotation"])
    delta = leaf(d["lbs_delta"]) if "lbs_delta" in d else None
    m, cov, _, _, _ = K.skin_and_covariance(xyz, d["vert_ids"].to(DEV), smpl.weights, ft, sc, rot, float(d["mod"]), delta)
    _close(m, d["world"])
    torch.autograd.backward([m, cov], [d["g_world"].to(DEV), d["g_cov"].to(DEV)])
    assert _rel(params["poses"].grad, d["d_poses"]) < 1e-4
    assert _rel(xyz.grad, d["d_xyz"]) < 1e-4
    if cr is not None:
        assert _rel(cr.grad, d["d_correct_Rs"]) < 1e-4


def test_joint_tables_regress_the_reference_joints():
    """J_template + J_shapedirs . betas == J_regressor . (v_template + shapedirs . betas) (the reference's order, fp64)."""
    body = synth.body()
    smpl = _smpl(body)
    jt, js = K.joint_tables(smpl)
    jt2, js2 = K.joint_tables(smpl)
    assert jt.data_ptr() == jt2.data_ptr() and js.data_ptr() == js2.data_ptr()         # cached per model
    _close(jt, body.J_regressor.double() @ body.verts.double(), 2e-6)
    _close(js, torch.einsum("jv,vck->jck", body.J_regressor.double(), body.shapedirs.double()), 2e-6)
    betas = synth.frame_params(4).params["shapes"].double().reshape(-1)
    ref = body.J_regressor.double() @ (body.verts.double() + (body.shapedirs.double() * betas).sum(-1))
    _close(jt.cpu().double() + (js.cpu().double() * betas).sum(-1), ref, 2e-6)


def test_argument_errors():
    smpl = _smpl(synth.body())
    fr = synth.frame_params(0)
    with pytest.raises(RuntimeError, match="72"):
        K.smpl_frame(smpl, torch.zeros(69, device=DEV), fr.params["shapes"].to(DEV))
    with pytest.raises(RuntimeError, match="23"):
        K.smpl_frame(smpl, fr.params["poses"].to(DEV), fr.params["shapes"].to(DEV), torch.zeros(22, 3, 3, device=DEV))
    with pytest.raises(RuntimeError, match="float32"):
        K.smpl_frame(smpl, fr.params["poses"].to(DEV).double(), fr.params["shapes"].to(DEV))
