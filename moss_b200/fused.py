"""Fused skin + rasterize: MOSS's one-view-per-iteration render (north_star subsystem 1).

What ``render()`` does from ``pc.get_xyz`` to the rasterizer's outputs (gaussian_renderer/__init__.py:57-125) --

    _, means3D, _, transforms, _ = pc.coarse_deform_c2source(xyz[None], smpl_param, big_pose_smpl_param, ...)
    cov3D_precomp = pc.get_covariance(scaling_modifier, transforms.squeeze())
    rendered_image, radii, depth, alpha = rasterizer(means3D=means3D, means2D=means2D, shs=shs,
                                                     colors_precomp=None, opacities=opacity, cov3D_precomp=cov3D_precomp)

-- as ONE call:

    ft = skinning.frame_tables(smpl, smpl_param, big_pose_smpl_param, correct_Rs, cache)
    image, radii, depth, alpha, transforms = skin_and_rasterize(raster_settings, xyz, vert_ids, smpl.weights, ft,
                                                                pc.get_scaling, pc._rotation, opacity, means2D, shs=shs)

The skinned centre and the transformed covariance live in registers of the per-Gaussian preprocess kernel
(``moss_skin_rasterize_forward`` / ``_backward``): 36 bytes per Gaussian that the two-call form writes and reads back, in
both directions, never touch HBM, and one launch disappears each way.  For many cameras of one frame keep the two-call
form (skinning once, ``rasterize_batch`` over the views).  No torch fallback, no CPU path.
"""
import ctypes as C

import torch

from . import _abi, rasterizer as R, skinning as K

_EMPTY = R._EMPTY


class _SkinRasterize(torch.autograd.Function):
    """differentiable: xyz, scaling, rotation, lbs_delta, A_tgt, poseoff_tgt, shs, colors_precomp, opacities, means2D."""

    @staticmethod
    def forward(ctx, xyz, scaling, rotation, lbs_delta, A_tgt, poseoff_tgt, shs, colors_precomp, opacities, means2D,
                vert_ids, weights, A_big, poseoff_big, shapeoff, Rg, Th, raster_settings, want_transforms):
        xyz = K._dev_f32(xyz, "xyz")
        if xyz.ndim != 2 or xyz.shape[1] != 3:
            raise RuntimeError("xyz must have dimensions (num_points, 3)")
        dev, P = xyz.device, xyz.shape[0]
        scaling, rotation = K._dev_f32(scaling, "scaling"), K._dev_f32(rotation, "rotation")
        if scaling is None or rotation is None:
            raise RuntimeError("the fused call builds the covariance: scaling and rotation are required")
        lbs_delta = K._dev_f32(lbs_delta, "lbs_delta")
        if lbs_delta is not None:
            lbs_delta = lbs_delta.reshape(P, 24)
        A_tgt, A_big = K._dev_f32(A_tgt, "A_tgt"), K._dev_f32(A_big, "A_big")
        poseoff_tgt, poseoff_big, shapeoff = (K._dev_f32(poseoff_tgt, "poseoff_tgt"), K._dev_f32(poseoff_big, "poseoff_big"),
                                              K._dev_f32(shapeoff, "shapeoff"))
        weights, Rg, Th = K._dev_f32(weights, "weights"), K._dev_f32(Rg, "R"), K._dev_f32(Th, "Th")
        if vert_ids.dtype != torch.int32:
            vert_ids = vert_ids.to(torch.int32)
        vert_ids = vert_ids.reshape(-1).contiguous()
        if not vert_ids.is_cuda or vert_ids.shape[0] != P:
            raise RuntimeError("vert_ids must be a CUDA tensor with one entry per Gaussian")
        rs = raster_settings
        k = K._fill_skin(xyz, vert_ids, weights, lbs_delta, A_big, A_tgt, poseoff_big, shapeoff, poseoff_tgt, Rg, Th,
                         scaling, rotation, float(rs.scale_modifier))
        with _abi.on_device(dev):
            transforms = torch.empty((P, 3, 3), dtype=torch.float32, device=dev) if want_transforms else None
        keep = (xyz, vert_ids, weights, lbs_delta, A_big, A_tgt, poseoff_big, shapeoff, poseoff_tgt, Rg, Th, scaling, rotation)
        sh = shs if shs is not None else _EMPTY
        cp = colors_precomp if colors_precomp is not None else _EMPTY
        color, radii, depth, alpha, st = R._forward(xyz, sh, cp, opacities, _EMPTY, _EMPTY, _EMPTY, [rs], R._default_flags, False,
                                                    skin=(k, keep, None, transforms))
        ctx.st, ctx.skin = st, k
        ctx.has = (lbs_delta is not None, poseoff_tgt is not None, shs is not None)
        ctx.V_smpl = weights.shape[0]
        ctx.save_for_backward(radii)
        outs = (color, radii, depth, alpha, transforms if want_transforms else torch.empty(0, device=dev))
        ctx.mark_non_differentiable(radii, outs[4])
        return outs

    @staticmethod
    def backward(ctx, g_color, _g_radii, g_depth, g_alpha, _g_tr):
        lib = _abi.load()
        (radii,) = ctx.saved_tensors
        st, k = ctx.st, ctx.skin
        sc = st.scene
        P, M, H, W = sc.P, sc.M, sc.H, sc.W
        dev = radii.device
        has_delta, has_pt, has_sh = ctx.has
        need = ctx.needs_input_grad
        with _abi.on_device(dev):
            def _g(t, shape, required=False):
                if t is None:
                    return torch.zeros(shape, dtype=torch.float32, device=dev) if required else None
                return R._f32c(t, "grad").expand(shape).contiguous() if tuple(t.shape) != tuple(shape) else R._f32c(t, "grad")
            g_color = _g(g_color, (3, H, W), required=True)
            g_depth, g_alpha = _g(g_depth, (1, H, W)), _g(g_alpha, (1, H, W))
            # one packed buffer, every row written by the kernels (nothing zero-filled): front-end gradients first
            n_sh = 3 * M * P if has_sh else 0
            sizes = [3 * P, 3 * P, 4 * P, 24 * P if (has_delta and need[3]) else 0, n_sh, P, 3 * P, 3 * P, 4 * P]
            offs, o = [], 0
            for n in sizes:
                offs.append(o); o += (n + 3) // 4 * 4
            buf = torch.empty((o,), dtype=torch.float32, device=dev)
            view = lambda i, *shape: buf[offs[i]:offs[i] + sizes[i]].view(*shape)  # noqa: E731
            d_xyz, d_scaling, d_rotation = view(0, P, 3), view(1, P, 3), view(2, P, 4)
            d_delta = view(3, P, 24) if sizes[3] else None
            d_sh = view(4, P, M, 3) if has_sh else None
            d_opacity, d_means2D, d_colors, d_conic = view(5, P, 1), view(6, P, 3), view(7, P, 3), view(8, P, 2, 2)
            d_A = torch.empty((24, 4, 4), dtype=torch.float32, device=dev) if need[4] else None
            d_po = torch.empty((ctx.V_smpl, 3), dtype=torch.float32, device=dev) if (has_pt and need[5]) else None
            p = K._p
            status = lib.moss_skin_rasterize_backward(
                _abi.stream_ptr(dev), C.byref(sc), C.byref(k), st.flags, p(radii),
                g_color.data_ptr(), p(g_depth), p(g_alpha), p(d_means2D), p(d_conic), p(d_opacity), p(d_colors), p(d_sh),
                p(d_xyz), p(d_scaling), p(d_rotation), p(d_delta), p(d_A), p(d_po))
        _abi.check(status, "moss_skin_rasterize_backward")
        ctx.st = None           # workspaces go back to the allocator
        return (d_xyz, d_scaling, d_rotation, d_delta, d_A, d_po, d_sh, (None if has_sh else d_colors), d_opacity, d_means2D) + (None,) * 9


def skin_and_rasterize(raster_settings, xyz, vert_ids, weights, ft: K.FrameTables, scaling, rotation, opacities,
                       means2D=None, shs=None, colors_precomp=None, lbs_delta=None, return_transforms=True):
    """-> (color [3,H,W], radii [P], depth [1,H,W], alpha [1,H,W], transforms [P,3,3] or None): the same values as
    ``skinning.skin_and_covariance`` followed by ``GaussianRasterizer(raster_settings)(means3D, means2D, opacities,
    shs/colors_precomp, cov3D_precomp=...)``, with gradients for xyz, scaling, rotation, lbs_delta, the frame tables
    (ft.A_tgt, ft.poseoff_tgt), shs / colors_precomp, opacities and ``means2D`` (the screen-space gradient sink
    ``viewspace_points``, gaussian_renderer/__init__.py:36-40)."""
    if (shs is None) == (colors_precomp is None):
        raise Exception('Please provide excatly one of either SHs or precomputed colors!')
    if means2D is None:
        means2D = torch.zeros_like(xyz, requires_grad=False)
    c, r, d, a, tr = _SkinRasterize.apply(xyz, scaling, rotation, lbs_delta, ft.A_tgt, ft.poseoff_tgt, shs, colors_precomp,
                                          opacities, means2D, vert_ids, weights, ft.A_big, ft.poseoff_big, ft.shapeoff,
                                          ft.R, ft.Th, raster_settings, bool(return_transforms))
    return c, r, d, a, (tr if return_transforms else None)
