"""Host-side mirror of MOSS's skinning + covariance front end, over the fused sm_100a kernel.

What render() does between ``pc.get_xyz`` and the rasterizer call (gaussian_renderer/__init__.py:57-85):

    _, means3D, bweights, transforms, translation = pc.coarse_deform_c2source(xyz[None], smpl_param,
                                                        big_pose_smpl_param, big_pose_world_vertex[None], ...)
    cov3D_precomp = pc.get_covariance(scaling_modifier, transforms.squeeze())

becomes

    ft = frame_tables(smpl, smpl_param, big_pose_smpl_param, correct_Rs)          # per frame, Gaussian-independent
    means3D, cov3D_precomp, transforms, bweights, translation = skin_and_covariance(
        xyz, vert_ids, smpl.weights, ft, pc.get_scaling, pc._rotation, scaling_modifier, lbs_delta)

``skin_and_covariance`` is ONE kernel forward and ONE backward through the C ABI (``moss_skin_forward`` /
``moss_skin_backward``, include/moss_splat.h); there is no torch fallback.  ``frame_tables`` is the 24-joint
kinematic chain and the blend-shape offset tables (scene/gaussian_model.py:838-892, 965-1064): ONE launch forward
and ONE backward as well (``moss_smpl_frame_forward`` / ``moss_smpl_frame_backward``, csrc/smpl_frame.cuh; SURVEY.md
section 8(f) rank 1); its big-pose half is constant over a run and is cached by ``FrameCache``.
"""
import ctypes as C
from typing import NamedTuple, Optional

import torch

from . import _abi


class Smpl(NamedTuple):
    """The entries of GaussianModel.SMPL_NEUTRAL the path reads (scene/gaussian_model.py:88-104), on the device."""
    v_template: torch.Tensor   # [V,3]
    shapedirs: torch.Tensor    # [V,3,10]
    posedirs: torch.Tensor     # [V,3,207]
    J_regressor: torch.Tensor  # [24,V]
    weights: torch.Tensor      # [V,24]
    parents: torch.Tensor      # [24] int64 (kintree_table[0])


class FrameTables(NamedTuple):
    A_big: torch.Tensor        # [24,4,4]
    A_tgt: torch.Tensor        # [24,4,4]
    poseoff_big: torch.Tensor  # [V,3]
    shapeoff: torch.Tensor     # [V,3]
    poseoff_tgt: torch.Tensor  # [V,3]
    R: torch.Tensor            # [3,3]
    Th: torch.Tensor           # [3]


def _dev_f32(t, name):
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (moss_b200 has no CPU path)")
    if t.dtype != torch.float32:
        raise RuntimeError(f"{name} must be float32")
    return t.contiguous()


def _p(t):
    return None if t is None else t.data_ptr()


_frame_scratch = {}     # (device index, stream handle) -> zero-initialised scratch of the frame kernels (self-cleaning)
_parents_i32 = {}       # (data_ptr, version) -> (source tensor kept alive, int32 copy)


def _scratch_for(dev, stream, nbytes):
    key = (dev.index, stream)
    t = _frame_scratch.get(key)
    if t is None or t.numel() < nbytes:
        t = torch.zeros((nbytes,), dtype=torch.uint8, device=dev)
        _frame_scratch[key] = t
    return t


def _parents32(parents):
    if parents.dtype == torch.int32 and parents.is_contiguous():
        return parents
    key = (parents.data_ptr(), parents._version)
    hit = _parents_i32.get(key)
    if hit is None or hit[0] is not parents:
        if len(_parents_i32) > 16:
            _parents_i32.clear()
        hit = (parents, parents.to(torch.int32).contiguous())
        _parents_i32[key] = hit
    return hit[1]


_joint_tables = {}      # (data_ptr, version) of the three source tables -> (sources kept alive, J_template, J_shapedirs)


def joint_tables(smpl):
    """J_template [24,3] = J_regressor . v_template and J_shapedirs [24,3,S] = J_regressor . shapedirs: constants of an SMPL
    model (``moss_smpl_joint_tables``, one launch, cached), so that the per-frame kernel regresses the joints
    (scene/gaussian_model.py:1007, :1021) without a pass over the vertices."""
    src = (smpl.v_template, smpl.shapedirs, smpl.J_regressor)
    key = tuple((t.data_ptr(), t._version) for t in src)
    hit = _joint_tables.get(key)
    if hit is not None and all(a is b for a, b in zip(hit[0], src)):
        return hit[1], hit[2]
    for name, t in zip(("v_template", "shapedirs", "J_regressor"), src):
        if not t.is_cuda:
            raise RuntimeError(f"smpl.{name} must be a CUDA tensor (moss_b200 has no CPU path)")
        if not (t.dtype == torch.float32 and t.is_contiguous() and t.device == src[0].device):
            raise RuntimeError(f"smpl.{name} must be a contiguous float32 CUDA tensor on one device")
    V, S = smpl.v_template.shape[0], smpl.shapedirs.shape[-1]
    if smpl.J_regressor.shape != (24, V) or smpl.shapedirs.shape[:2] != (V, 3):
        raise RuntimeError("SMPL table shapes do not match (v_template [V,3], shapedirs [V,3,S], J_regressor [24,V])")
    dev = smpl.v_template.device
    with _abi.on_device(dev):
        jt = torch.empty((24, 3), dtype=torch.float32, device=dev)
        js = torch.empty((24, 3, S), dtype=torch.float32, device=dev)
        st = _abi.load().moss_smpl_joint_tables(_abi.stream_ptr(dev), V, S, _p(smpl.v_template),
                                                _p(smpl.shapedirs), _p(smpl.J_regressor), _p(jt), _p(js))
    _abi.check(st, "moss_smpl_joint_tables")
    torch.cuda.current_stream(dev).synchronize()      # once per model: later frames may run on any stream
    if len(_joint_tables) > 8:
        _joint_tables.clear()
    _joint_tables[key] = (src, jt, js)
    return jt, js


def _fill_frame(smpl, poses, shapes, correct_Rs, scratch, parents, tables=None):
    jt, js = tables if tables is not None else joint_tables(smpl)
    f = _abi.MossSmplFrame()
    f.V, f.n_shapes, f.shape_stride = smpl.v_template.shape[0], shapes.numel(), smpl.shapedirs.shape[-1]
    f.shapedirs, f.posedirs, f.J_template, f.J_shapedirs = _p(smpl.shapedirs), _p(smpl.posedirs), _p(jt), _p(js)
    f.parents, f.poses, f.shapes, f.correct_Rs = _p(parents), _p(poses), _p(shapes), _p(correct_Rs)
    f.scratch, f.scratch_bytes = _p(scratch), (0 if scratch is None else scratch.numel())
    return f


class _SmplFrame(torch.autograd.Function):
    """poses [72] (and correct_Rs [23,3,3]) -> A [24,4,4], pose_offsets [V,3], shape_offset [V,3]: ONE launch forward
    (``moss_smpl_frame_forward``) and ONE backward (``moss_smpl_frame_backward``); shapes and the SMPL tables are
    constants of a run.  Replaces scene/gaussian_model.py:855-861, 870-871, 878-893, 965-1064 for one frame."""

    @staticmethod
    def forward(ctx, poses, correct_Rs, shapes, smpl):
        lib = _abi.load()
        ctx.in_shapes = (tuple(poses.shape), None if correct_Rs is None else tuple(correct_Rs.shape))
        poses = _dev_f32(poses.reshape(-1), "poses")
        if poses.numel() != 72:
            raise RuntimeError("poses must hold 24 axis-angle vectors (72 floats)")
        dev = poses.device
        shapes = _dev_f32(shapes.reshape(-1), "shapes")
        if correct_Rs is not None:
            correct_Rs = _dev_f32(correct_Rs.reshape(-1), "correct_Rs")
            if correct_Rs.numel() != 23 * 9:
                raise RuntimeError("correct_Rs must hold 23 3x3 matrices")
        tables = joint_tables(smpl)             # validates v_template / shapedirs / J_regressor
        t = smpl.posedirs
        if not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous() and t.device == dev):
            raise RuntimeError("smpl.posedirs must be a contiguous float32 CUDA tensor on the device of poses")
        V = smpl.v_template.shape[0]
        if shapes.numel() > smpl.shapedirs.shape[-1] or smpl.posedirs.numel() != V * 3 * 207:
            raise RuntimeError("SMPL table shapes do not match (shapedirs [V,3,S] with S >= len(shapes), posedirs [V,3,207])")
        parents = _parents32(smpl.parents)
        with _abi.on_device(dev):
            stream = _abi.stream_ptr(dev)
            out = torch.empty((24 * 16 + 24 * 3 + 6 * V,), dtype=torch.float32, device=dev)
            A, joints = out[:384].view(24, 4, 4), out[384:456]
            poseoff, shapeoff = out[456:456 + 3 * V].view(V, 3), out[456 + 3 * V:].view(V, 3)
            f = _fill_frame(smpl, poses, shapes, correct_Rs, None, parents, tables)
            st = lib.moss_smpl_frame_forward(stream, C.byref(f), _p(A), None, _p(joints), _p(poseoff), _p(shapeoff))
        _abi.check(st, "moss_smpl_frame_forward")
        ctx.smpl = smpl
        ctx.has_cr = correct_Rs is not None
        ctx.save_for_backward(poses, shapes, A, joints, *([correct_Rs] if correct_Rs is not None else []))
        ctx.mark_non_differentiable(shapeoff)
        return A, poseoff, shapeoff

    @staticmethod
    def backward(ctx, d_A, d_poseoff, _ds):
        lib = _abi.load()
        sv = ctx.saved_tensors
        poses, shapes, A, joints = sv[:4]
        correct_Rs = sv[4] if ctx.has_cr else None
        smpl, dev = ctx.smpl, poses.device
        V = smpl.v_template.shape[0]
        d_A = _dev_f32(d_A, "grad A") if d_A is not None else None
        d_poseoff = _dev_f32(d_poseoff, "grad pose_offsets") if d_poseoff is not None else None
        with _abi.on_device(dev):
            stream = _abi.stream_ptr(dev)
            scratch = _scratch_for(dev, stream, lib.moss_smpl_frame_scratch_bytes(V))
            d_poses = torch.empty((72,), dtype=torch.float32, device=dev)
            d_cr = torch.empty((23, 3, 3), dtype=torch.float32, device=dev) if (ctx.has_cr and ctx.needs_input_grad[1]) else None
            f = _fill_frame(smpl, poses, shapes, correct_Rs, scratch, _parents32(smpl.parents))
            st = lib.moss_smpl_frame_backward(stream, C.byref(f), _p(A), _p(joints), _p(d_A), _p(d_poseoff), _p(d_poses), _p(d_cr))
        _abi.check(st, "moss_smpl_frame_backward")
        return d_poses.view(ctx.in_shapes[0]), (None if d_cr is None else d_cr.view(ctx.in_shapes[1])), None, None


def smpl_frame(smpl: Smpl, poses, shapes, correct_Rs=None):
    """-> (A [24,4,4], pose_offsets [V,3], shape_offset [V,3]) of one frame; differentiable w.r.t. ``poses`` and
    ``correct_Rs``.  The values get_transform_params_torch(...)[0] (scene/gaussian_model.py:998-1031) and the two
    blend-shape tables of coarse_deform_c2source (:855-861 / :878-893, :870-871) hold before the per-Gaussian gather."""
    return _SmplFrame.apply(poses, correct_Rs, shapes, smpl)


class FrameCache:
    """The big-pose half of the tables never changes during a run (scene/dataset_readers.py:583-590)."""

    def __init__(self):
        self.key = None
        self.val = None

    def get(self, smpl, t_params):
        # keyed on storage address AND in-place version of the three tensors; the tensors themselves are kept
        # referenced while cached, so the caching allocator cannot hand their address to different contents
        src = (t_params["poses"], t_params["shapes"], smpl.v_template)
        key = tuple((t.data_ptr(), t._version, tuple(t.shape)) for t in src)
        if self.key != key:
            with torch.no_grad():
                A, po, _ = smpl_frame(smpl, t_params["poses"], t_params["shapes"])
            self.val = (A, po)
            self.key = key
            self.src = src
        return self.val


def frame_tables(smpl: Smpl, params, t_params, correct_Rs=None, cache: Optional[FrameCache] = None) -> FrameTables:
    """Per-frame half of coarse_deform_c2source (scene/gaussian_model.py:838-892).  ``params`` / ``t_params`` are the
    reference's ``smpl_param`` / ``big_pose_smpl_param`` dicts (poses [1,72], shapes [1,10], R [3,3], Th [1,3]).
    Two launches without a cache (big pose, target pose), one with it; gradients flow to ``params['poses']`` and
    ``correct_Rs`` through the native backward."""
    if cache is not None:
        A_big, po_big = cache.get(smpl, t_params)
    else:
        with torch.no_grad():
            A_big, po_big, _ = smpl_frame(smpl, t_params["poses"], t_params["shapes"])
    A, po, so = smpl_frame(smpl, params["poses"], params["shapes"], correct_Rs)
    return FrameTables(A_big, A, po_big, so, po, params["R"].reshape(3, 3), params["Th"].reshape(-1))


def _fill_skin(xyz, vert_ids, weights, lbs_delta, A_big, A_tgt, po_big, so, po_tgt, R, Th, scaling, rotation, mod):
    k = _abi.MossSkin()
    k.P, k.V = xyz.shape[0], weights.shape[0]
    k.xyz, k.vert_ids, k.weights, k.lbs_delta = _p(xyz), _p(vert_ids), _p(weights), _p(lbs_delta)
    k.A_big, k.A_tgt = _p(A_big), _p(A_tgt)
    k.poseoff_big, k.shapeoff, k.poseoff_tgt = _p(po_big), _p(so), _p(po_tgt)
    k.Rg, k.Th, k.scaling, k.rotation = _p(R), _p(Th), _p(scaling), _p(rotation)
    k.scale_modifier = float(mod)
    return k


class _SkinCovariance(torch.autograd.Function):
    """inputs: xyz, scaling, rotation, lbs_delta, A_tgt, poseoff_tgt are differentiable; the rest are constants."""

    @staticmethod
    def forward(ctx, xyz, scaling, rotation, lbs_delta, A_tgt, poseoff_tgt, vert_ids, weights, A_big, poseoff_big,
                shapeoff, R, Th, scaling_modifier, return_transl):
        lib = _abi.load()
        xyz = _dev_f32(xyz, "xyz")
        if xyz.ndim != 2 or xyz.shape[1] != 3:
            raise RuntimeError("xyz must have dimensions (num_points, 3)")
        dev = xyz.device
        P = xyz.shape[0]
        scaling, rotation = _dev_f32(scaling, "scaling"), _dev_f32(rotation, "rotation")
        lbs_delta = _dev_f32(lbs_delta, "lbs_delta")
        if lbs_delta is not None:
            lbs_delta = lbs_delta.reshape(P, 24)
        A_tgt, A_big = _dev_f32(A_tgt, "A_tgt"), _dev_f32(A_big, "A_big")
        poseoff_tgt, poseoff_big, shapeoff = (_dev_f32(poseoff_tgt, "poseoff_tgt"), _dev_f32(poseoff_big, "poseoff_big"),
                                              _dev_f32(shapeoff, "shapeoff"))
        weights, R, Th = _dev_f32(weights, "weights"), _dev_f32(R, "R"), _dev_f32(Th, "Th")
        if vert_ids.dtype != torch.int32:
            vert_ids = vert_ids.to(torch.int32)
        vert_ids = vert_ids.reshape(-1).contiguous()
        if not vert_ids.is_cuda or vert_ids.shape[0] != P:
            raise RuntimeError("vert_ids must be a CUDA tensor with one entry per Gaussian")
        with _abi.on_device(dev):
            means3D = torch.empty((P, 3), dtype=torch.float32, device=dev)
            has_cov = scaling is not None and rotation is not None
            cov3D = torch.empty((P, 6), dtype=torch.float32, device=dev) if has_cov else None
            transforms = torch.empty((P, 3, 3), dtype=torch.float32, device=dev)
            bweights = torch.empty((P, 24), dtype=torch.float32, device=dev)
            translation = torch.empty((P, 3), dtype=torch.float32, device=dev) if return_transl else None
            k = _fill_skin(xyz, vert_ids, weights, lbs_delta, A_big, A_tgt, poseoff_big, shapeoff, poseoff_tgt, R, Th,
                           scaling, rotation, scaling_modifier)
            st = lib.moss_skin_forward(_abi.stream_ptr(dev), C.byref(k), _p(means3D), _p(cov3D),
                                       _p(transforms), _p(bweights), _p(translation))
        _abi.check(st, "moss_skin_forward")
        ctx.mod = float(scaling_modifier)
        ctx.has = (scaling is not None, lbs_delta is not None, poseoff_big is not None, shapeoff is not None,
                   poseoff_tgt is not None)
        saved = [xyz, vert_ids, weights, A_big, A_tgt, R, Th]
        for t in (scaling, rotation, lbs_delta, poseoff_big, shapeoff, poseoff_tgt):
            if t is not None:
                saved.append(t)
        ctx.save_for_backward(*saved)
        outs = (means3D, cov3D if has_cov else torch.empty(0, device=dev), transforms, bweights,
                translation if return_transl else torch.empty(0, device=dev))
        ctx.mark_non_differentiable(outs[2], outs[3], outs[4])
        return outs

    @staticmethod
    def backward(ctx, g_means3D, g_cov3D, _gt, _gb, _gtr):
        lib = _abi.load()
        sv = list(ctx.saved_tensors)
        xyz, vert_ids, weights, A_big, A_tgt, R, Th = sv[:7]
        rest = sv[7:]
        has_cov, has_delta, has_pb, has_so, has_pt = ctx.has
        scaling = rotation = lbs_delta = po_big = so = po_tgt = None
        if has_cov:
            scaling, rotation = rest[0], rest[1]; rest = rest[2:]
        if has_delta:
            lbs_delta = rest[0]; rest = rest[1:]
        if has_pb:
            po_big = rest[0]; rest = rest[1:]
        if has_so:
            so = rest[0]; rest = rest[1:]
        if has_pt:
            po_tgt = rest[0]; rest = rest[1:]
        dev = xyz.device
        P, V = xyz.shape[0], weights.shape[0]
        need = ctx.needs_input_grad        # xyz, scaling, rotation, lbs_delta, A_tgt, poseoff_tgt
        with _abi.on_device(dev):
            g_means3D = (torch.zeros((P, 3), dtype=torch.float32, device=dev) if g_means3D is None
                         else _dev_f32(g_means3D, "grad means3D"))
            g_cov3D = _dev_f32(g_cov3D, "grad cov3D") if (has_cov and g_cov3D is not None and g_cov3D.numel()) else None
            d_xyz = torch.empty((P, 3), dtype=torch.float32, device=dev)
            d_scaling = torch.empty((P, 3), dtype=torch.float32, device=dev) if has_cov else None
            d_rotation = torch.empty((P, 4), dtype=torch.float32, device=dev) if has_cov else None
            d_delta = torch.empty((P, 24), dtype=torch.float32, device=dev) if (has_delta and need[3]) else None
            d_A = torch.empty((24, 4, 4), dtype=torch.float32, device=dev) if need[4] else None
            d_po = torch.empty((V, 3), dtype=torch.float32, device=dev) if (has_pt and need[5]) else None
            k = _fill_skin(xyz, vert_ids, weights, lbs_delta, A_big, A_tgt, po_big, so, po_tgt, R, Th, scaling, rotation,
                           ctx.mod)
            st = lib.moss_skin_backward(_abi.stream_ptr(dev), C.byref(k), _p(g_means3D), _p(g_cov3D),
                                        _p(d_xyz), _p(d_scaling), _p(d_rotation), _p(d_delta), _p(d_A), _p(d_po))
        _abi.check(st, "moss_skin_backward")
        return (d_xyz, d_scaling, d_rotation, d_delta, d_A, d_po) + (None,) * 9


def skin_and_covariance(xyz, vert_ids, weights, ft: FrameTables, scaling=None, rotation=None, scaling_modifier=1.0,
                        lbs_delta=None, return_transl=False):
    """-> (means3D [P,3], cov3D_precomp [P,6] or None, transforms [P,3,3], bweights [P,24], translation [P,3] or None).

    Same values as ``coarse_deform_c2source(...)[1:]`` + ``get_covariance(scaling_modifier, transforms)``
    (scene/gaussian_model.py:820-923, 37-44).  ``vert_ids`` is the nearest-vertex lookup the reference recomputes with
    knn_cuda every call (:827); ``lbs_delta`` is its ``lbs_weights`` argument."""
    m, c, tr, bw, tl = _SkinCovariance.apply(xyz, scaling, rotation, lbs_delta, ft.A_tgt, ft.poseoff_tgt, vert_ids,
                                             weights, ft.A_big, ft.poseoff_big, ft.shapeoff, ft.R, ft.Th,
                                             scaling_modifier, return_transl)
    return m, (c if c.numel() else None), tr, bw, (tl if return_transl else None)


class VertexGrid:
    """A uniform grid over a vertex set (``moss_vertex_grid_build``): build once for the big-pose world vertices -- they are
    constants of a run (scene/dataset_readers.py:586-601) -- and every later nearest-vertex query examines a few dozen
    vertices instead of all of them, with answers identical to the exhaustive search."""

    def __init__(self, t_vertices):
        lib = _abi.load()
        v = _dev_f32(t_vertices.detach().reshape(-1, 3), "t_vertices")
        if v.shape[0] < 1:
            raise RuntimeError("t_vertices is empty")
        self.V, self.device = v.shape[0], v.device
        nbytes = lib.moss_vertex_grid_bytes(self.V)
        with _abi.on_device(v.device):
            self.buf = torch.empty((nbytes,), dtype=torch.uint8, device=v.device)
            st = lib.moss_vertex_grid_build(_abi.stream_ptr(v.device), self.V, _p(v), _p(self.buf), nbytes)
        _abi.check(st, "moss_vertex_grid_build")
        self.src = t_vertices            # keeps the identity the cache is keyed on alive


_vertex_grids = {}      # (data_ptr, version, shape) of a vertex tensor -> VertexGrid


def _grid_for(t_vertices):
    key = (t_vertices.data_ptr(), t_vertices._version, tuple(t_vertices.shape))
    g = _vertex_grids.get(key)
    if g is None or g.src is not t_vertices:
        if len(_vertex_grids) > 8:
            _vertex_grids.clear()
        g = _vertex_grids[key] = VertexGrid(t_vertices)
    return g


def nearest_vertex(xyz, t_vertices, return_distance=False, grid="auto"):
    """int32 [P]: index of the nearest SMPL vertex of every Gaussian -- what the reference asks knn_cuda for on every
    render, ``_, vert_ids = self.knn(smpl_pts.float(), query_pts.float())`` with ``KNN(k=1, transpose_mode=True)``
    (scene/gaussian_model.py:85, :827).  ``xyz`` [P,3] or [1,P,3], ``t_vertices`` [V,3] or [1,V,3] (the big-pose world
    vertices).  With ``return_distance`` also the Euclidean distance [P] (``distance, _ = self.knn(...)``, :657).
    Gaussians move little between densifications, so a training loop may keep the result and refresh it only when the
    set of Gaussians changes.
    ``grid``: "auto" (default) searches through a ``VertexGrid`` of ``t_vertices``, built on first use and cached while the
    same tensor object is passed unmodified; a ``VertexGrid`` instance uses that one; None runs the exhaustive kernel.
    All three give the same answer."""
    lib = _abi.load()
    q = _dev_f32(xyz.detach().reshape(-1, 3), "xyz")
    v = _dev_f32(t_vertices.detach().reshape(-1, 3), "t_vertices")
    if v.device != q.device:
        raise RuntimeError("xyz and t_vertices must be on the same device")
    P, V = q.shape[0], v.shape[0]
    if V < 1:
        raise RuntimeError("t_vertices is empty")
    ids = torch.empty((P,), dtype=torch.int32, device=q.device)
    dist = torch.empty((P,), dtype=torch.float32, device=q.device) if return_distance else None
    if isinstance(grid, str) and grid == "auto":
        grid = _grid_for(t_vertices) if V >= 64 else None
    with _abi.on_device(q.device):
        stream = _abi.stream_ptr(q.device)
        if grid is not None:
            if grid.V != V or grid.device != q.device:
                raise RuntimeError("the VertexGrid was built for another vertex set")
            status = lib.moss_nearest_vertex_grid(stream, P, _p(q), V, _p(grid.buf), grid.buf.numel(), _p(ids), _p(dist))
        else:
            status = lib.moss_nearest_vertex(stream, P, _p(q), V, _p(v), _p(ids), _p(dist))
    _abi.check(status, "moss_nearest_vertex")
    return (ids, dist) if return_distance else ids
