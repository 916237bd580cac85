"""ctypes binding of the C-ABI library (include/moss_splat.h).

The product path has NO fallback: if ``libmoss_splat.so`` is missing or does not export the
declared symbols, importing/using the rasterizer raises.  Build it with ``python -m moss_b200.build``
(or ``__graft_entry__.build()``).
"""
import contextlib
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MOSS_SPLAT_LIB") or os.path.join(_HERE, "_lib", "libmoss_splat.so")    # override: A/B builds of the same ABI

MOSS_OK = 0
MOSS_ERR_BINNING_OVERFLOW = 1
MOSS_ERR_INVALID_ARGUMENT = 2
MOSS_ERR_WORKSPACE_TOO_SMALL = 3
MOSS_ERR_CUDA = 4
MOSS_ERR_STATIC_SCRATCH_LIMIT = 5

FLAG_ASYNC = 1
FLAG_REBIN = 2
FLAG_EXPORT_KEYS = 4
FLAG_DEBUG = 8
FLAG_DETERMINISTIC = 16
FLAG_NO_CULL = 32
FLAG_SPLIT_SHIFT = 8          # bits 8-9: 0 auto, 1 whole-tile blocks, 2 half-tile, 3 quarter-tile

STATE = dict(depths=0, clamped=1, means2D=2, cov3D=3, conic_opacity=4, rgb=5, tiles_touched=7,
             keys_sorted=9, point_list=11, ranges=12, n_contrib=13, final_T=14, counters=15, view_base=16)
MAX_VIEWS = 32

# every symbol include/moss_splat.h declares
SYMBOLS = [
    "moss_geom_workspace_bytes", "moss_image_workspace_bytes", "moss_binning_workspace_bytes",
    "moss_binning_workspace_bytes_ex", "moss_binning_capacity", "moss_ckpt_workspace_bytes", "moss_rasterize_forward",
    "moss_rasterize_backward", "moss_mark_visible", "moss_export_state", "moss_abi_version",
    "moss_error_string", "moss_last_cuda_error", "moss_sizeof_scene", "moss_kernel_launches",
    "moss_profile_enable", "moss_profile_read", "moss_stage_name", "moss_nvtx_enable",
    "moss_geom_workspace_bytes_batch", "moss_image_workspace_bytes_batch", "moss_ckpt_workspace_bytes_batch",
    "moss_batch_grad_workspace_bytes", "moss_rasterize_forward_batch", "moss_rasterize_backward_batch", "moss_sizeof_view",
    "moss_export_state_batch", "moss_push_slice_floats", "moss_sizeof_push", "moss_rasterize_backward_push", "moss_push_allreduce",
    "moss_skin_forward", "moss_skin_backward", "moss_sizeof_skin", "moss_nearest_vertex", "moss_vertex_grid_bytes", "moss_vertex_grid_build", "moss_nearest_vertex_grid", "moss_densify_stats", "moss_adamw_step", "moss_sizeof_adam_group", "moss_peer_signal_bytes", "moss_peer_allreduce",
    "moss_smpl_frame_scratch_bytes", "moss_sizeof_smpl_frame", "moss_smpl_frame_forward", "moss_smpl_frame_backward", "moss_smpl_joint_tables", "moss_skin_rasterize_forward", "moss_skin_rasterize_backward",
    "moss_image_losses_scratch_bytes", "moss_sizeof_image_loss", "moss_image_losses", "moss_image_losses_static", "moss_mask_bounds", "moss_scale_gradients",
]
STAGE_COUNT = 19


class MossScene(C.Structure):
    _fields_ = [
        ("P", C.c_int), ("D", C.c_int), ("M", C.c_int), ("W", C.c_int), ("H", C.c_int),
        ("bg", C.c_void_p), ("means3D", C.c_void_p), ("shs", C.c_void_p), ("colors_precomp", C.c_void_p),
        ("opacities", C.c_void_p), ("scales", C.c_void_p), ("rotations", C.c_void_p),
        ("cov3D_precomp", C.c_void_p), ("viewmatrix", C.c_void_p), ("projmatrix", C.c_void_p),
        ("campos", C.c_void_p),
        ("scale_modifier", C.c_float), ("tanfovx", C.c_float), ("tanfovy", C.c_float),
        ("prefiltered", C.c_int),
        ("geom_ws", C.c_void_p), ("geom_bytes", C.c_size_t),
        ("image_ws", C.c_void_p), ("image_bytes", C.c_size_t),
        ("binning_ws", C.c_void_p), ("binning_bytes", C.c_size_t),
        ("ckpt_ws", C.c_void_p), ("ckpt_bytes", C.c_size_t),
    ]


class MossView(C.Structure):
    _fields_ = [("viewmatrix", C.c_void_p), ("projmatrix", C.c_void_p), ("campos", C.c_void_p), ("bg", C.c_void_p),
                ("tanfovx", C.c_float), ("tanfovy", C.c_float)]


class MossPush(C.Structure):
    _fields_ = [("rank", C.c_int), ("world", C.c_int), ("grad_base", C.c_void_p), ("numel", C.c_size_t),
                ("grad_ptrs", C.POINTER(C.c_void_p)), ("stage_ptrs", C.POINTER(C.c_void_p)),
                ("signal_ptrs", C.POINTER(C.c_void_p)), ("signal_pad_bytes", C.c_size_t),
                ("multicast_grad", C.c_void_p), ("error_word_dev", C.c_void_p)]


class MossAdamGroup(C.Structure):
    _fields_ = [("param", C.c_void_p), ("grad", C.c_void_p), ("exp_avg", C.c_void_p), ("exp_avg_sq", C.c_void_p),
                ("numel", C.c_size_t), ("lr", C.c_double), ("step", C.c_int)]


class MossSkin(C.Structure):
    _fields_ = [
        ("P", C.c_int), ("V", C.c_int),
        ("xyz", C.c_void_p), ("vert_ids", C.c_void_p), ("weights", C.c_void_p), ("lbs_delta", C.c_void_p),
        ("A_big", C.c_void_p), ("A_tgt", C.c_void_p), ("poseoff_big", C.c_void_p), ("shapeoff", C.c_void_p),
        ("poseoff_tgt", C.c_void_p), ("Rg", C.c_void_p), ("Th", C.c_void_p), ("scaling", C.c_void_p),
        ("rotation", C.c_void_p), ("scale_modifier", C.c_float),
    ]


class MossSmplFrame(C.Structure):
    _fields_ = [
        ("V", C.c_int), ("n_shapes", C.c_int), ("shape_stride", C.c_int),
        ("shapedirs", C.c_void_p), ("posedirs", C.c_void_p), ("J_template", C.c_void_p), ("J_shapedirs", C.c_void_p),
        ("parents", C.c_void_p), ("poses", C.c_void_p), ("shapes", C.c_void_p), ("correct_Rs", C.c_void_p),
        ("scratch", C.c_void_p), ("scratch_bytes", C.c_size_t),
    ]


class MossImageLoss(C.Structure):
    _fields_ = [
        ("W", C.c_int), ("H", C.c_int),
        ("image", C.c_void_p), ("alpha", C.c_void_p), ("gt_image", C.c_void_p), ("bkgd_mask", C.c_void_p),
        ("bound_mask", C.c_void_p), ("rect", C.c_void_p),
        ("w_l1", C.c_float), ("w_mask", C.c_float), ("w_ssim", C.c_float), ("w_s3im", C.c_float),
        ("scratch", C.c_void_p), ("scratch_bytes", C.c_size_t),
    ]


_lib = None


def load():
    """Load the shared library once; raise loudly when it is absent (no CPU / eager fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"moss_b200: native library {LIB_PATH} not built. Run `python -m moss_b200.build` "
            "(needs nvcc; cross-compiles for sm_100a without a GPU). There is no fallback path.")
    lib = C.CDLL(LIB_PATH)
    missing = [s for s in SYMBOLS if not hasattr(lib, s)]
    if missing:
        raise RuntimeError(f"moss_b200: {LIB_PATH} lacks symbols {missing}; rebuild it")
    vp = C.c_void_p
    lib.moss_geom_workspace_bytes.restype = C.c_size_t
    lib.moss_geom_workspace_bytes.argtypes = [C.c_int]
    lib.moss_image_workspace_bytes.restype = C.c_size_t
    lib.moss_image_workspace_bytes.argtypes = [C.c_int, C.c_int]
    lib.moss_binning_workspace_bytes.restype = C.c_size_t
    lib.moss_binning_workspace_bytes.argtypes = [C.c_longlong]
    lib.moss_binning_workspace_bytes_ex.restype = C.c_size_t
    lib.moss_binning_workspace_bytes_ex.argtypes = [C.c_longlong, C.c_int]
    lib.moss_ckpt_workspace_bytes.restype = C.c_size_t
    lib.moss_ckpt_workspace_bytes.argtypes = [C.c_longlong, C.c_int, C.c_int]
    lib.moss_binning_capacity.restype = C.c_longlong
    lib.moss_binning_capacity.argtypes = [C.c_size_t, C.c_int]
    lib.moss_rasterize_forward.restype = C.c_int
    lib.moss_rasterize_forward.argtypes = [vp, C.POINTER(MossScene), C.c_uint, vp, vp, vp, vp, C.POINTER(C.c_int)]
    lib.moss_rasterize_backward.restype = C.c_int
    lib.moss_rasterize_backward.argtypes = [vp, C.POINTER(MossScene), C.c_uint, vp] + [vp] * 12
    lib.moss_mark_visible.restype = C.c_int
    lib.moss_mark_visible.argtypes = [vp, C.c_int, vp, vp, vp, vp]
    lib.moss_export_state.restype = C.c_longlong
    lib.moss_export_state.argtypes = [vp, C.POINTER(MossScene), C.c_int, vp, C.c_longlong]
    lib.moss_abi_version.restype = C.c_int
    lib.moss_error_string.restype = C.c_char_p
    lib.moss_error_string.argtypes = [C.c_int]
    lib.moss_last_cuda_error.restype = C.c_int
    lib.moss_sizeof_scene.restype = C.c_size_t
    lib.moss_kernel_launches.restype = C.c_ulonglong
    lib.moss_profile_enable.restype = C.c_int
    lib.moss_profile_enable.argtypes = [C.c_int]
    lib.moss_nvtx_enable.restype = C.c_int
    lib.moss_nvtx_enable.argtypes = [C.c_int]
    lib.moss_profile_read.restype = C.c_int
    lib.moss_profile_read.argtypes = [C.POINTER(C.c_double), C.POINTER(C.c_ulonglong)]
    lib.moss_stage_name.restype = C.c_char_p
    lib.moss_stage_name.argtypes = [C.c_int]
    lib.moss_peer_signal_bytes.restype = C.c_size_t
    lib.moss_peer_signal_bytes.argtypes = [C.c_int]
    lib.moss_peer_allreduce.restype = C.c_int
    lib.moss_peer_allreduce.argtypes = [vp, C.c_int, C.c_int, vp, vp, C.c_size_t, C.c_size_t, C.c_size_t, C.c_uint, vp, vp]
    lib.moss_sizeof_skin.restype = C.c_size_t
    lib.moss_geom_workspace_bytes_batch.restype = C.c_size_t
    lib.moss_geom_workspace_bytes_batch.argtypes = [C.c_int, C.c_int]
    lib.moss_image_workspace_bytes_batch.restype = C.c_size_t
    lib.moss_image_workspace_bytes_batch.argtypes = [C.c_int, C.c_int, C.c_int]
    lib.moss_ckpt_workspace_bytes_batch.restype = C.c_size_t
    lib.moss_ckpt_workspace_bytes_batch.argtypes = [C.c_longlong, C.c_int, C.c_int, C.c_int]
    lib.moss_batch_grad_workspace_bytes.restype = C.c_size_t
    lib.moss_batch_grad_workspace_bytes.argtypes = [C.c_int, C.c_int, C.c_int]
    lib.moss_rasterize_forward_batch.restype = C.c_int
    lib.moss_rasterize_forward_batch.argtypes = [vp, C.POINTER(MossScene), C.c_int, C.POINTER(MossView), C.c_uint, vp, vp, vp, vp,
                                                 C.POINTER(C.c_int)]
    lib.moss_rasterize_backward_batch.restype = C.c_int
    lib.moss_rasterize_backward_batch.argtypes = [vp, C.POINTER(MossScene), C.c_int, C.POINTER(MossView), C.c_uint, vp] + [vp] * 12 + [vp, C.c_size_t]
    lib.moss_export_state_batch.restype = C.c_longlong
    lib.moss_export_state_batch.argtypes = [vp, C.POINTER(MossScene), C.c_int, C.c_int, vp, C.c_longlong]
    lib.moss_sizeof_view.restype = C.c_size_t
    lib.moss_push_slice_floats.restype = C.c_size_t
    lib.moss_push_slice_floats.argtypes = [C.c_size_t, C.c_int]
    lib.moss_sizeof_push.restype = C.c_size_t
    lib.moss_push_allreduce.restype = C.c_int
    lib.moss_push_allreduce.argtypes = [vp, C.POINTER(MossPush)]
    lib.moss_rasterize_backward_push.restype = C.c_int
    lib.moss_rasterize_backward_push.argtypes = [vp, C.POINTER(MossScene), C.c_int, C.POINTER(MossView), C.c_uint, vp] + [vp] * 12 + \
        [vp, C.c_size_t, C.POINTER(MossPush)]
    if lib.moss_sizeof_push() != C.sizeof(MossPush):
        raise RuntimeError("moss_b200: MossPush layout mismatch between _abi.py and the library; rebuild")
    if lib.moss_sizeof_view() != C.sizeof(MossView):
        raise RuntimeError("moss_b200: MossView layout mismatch between _abi.py and the library; rebuild")
    lib.moss_adamw_step.restype = C.c_int
    lib.moss_adamw_step.argtypes = [vp, C.c_int, C.POINTER(MossAdamGroup), C.c_double, C.c_double, C.c_double, C.c_double]
    lib.moss_sizeof_adam_group.restype = C.c_size_t
    lib.moss_densify_stats.restype = C.c_int
    lib.moss_densify_stats.argtypes = [vp, C.c_int, vp, vp, vp, vp, vp]
    lib.moss_nearest_vertex.restype = C.c_int
    lib.moss_nearest_vertex.argtypes = [vp, C.c_int, vp, C.c_int, vp, vp, vp]
    lib.moss_vertex_grid_bytes.restype = C.c_size_t
    lib.moss_vertex_grid_bytes.argtypes = [C.c_int]
    lib.moss_vertex_grid_build.restype = C.c_int
    lib.moss_vertex_grid_build.argtypes = [vp, C.c_int, vp, vp, C.c_size_t]
    lib.moss_nearest_vertex_grid.restype = C.c_int
    lib.moss_nearest_vertex_grid.argtypes = [vp, C.c_int, vp, C.c_int, vp, C.c_size_t, vp, vp]
    lib.moss_skin_forward.restype = C.c_int
    lib.moss_skin_forward.argtypes = [vp, C.POINTER(MossSkin)] + [vp] * 5
    lib.moss_skin_backward.restype = C.c_int
    lib.moss_skin_backward.argtypes = [vp, C.POINTER(MossSkin)] + [vp] * 8
    lib.moss_smpl_frame_scratch_bytes.restype = C.c_size_t
    lib.moss_smpl_frame_scratch_bytes.argtypes = [C.c_int]
    lib.moss_sizeof_smpl_frame.restype = C.c_size_t
    lib.moss_smpl_frame_forward.restype = C.c_int
    lib.moss_smpl_frame_forward.argtypes = [vp, C.POINTER(MossSmplFrame)] + [vp] * 5
    lib.moss_skin_rasterize_forward.restype = C.c_int
    lib.moss_skin_rasterize_forward.argtypes = [vp, C.POINTER(MossScene), C.POINTER(MossSkin), C.c_uint, vp, vp, vp, vp, C.POINTER(C.c_int), vp, vp]
    lib.moss_skin_rasterize_backward.restype = C.c_int
    lib.moss_skin_rasterize_backward.argtypes = [vp, C.POINTER(MossScene), C.POINTER(MossSkin), C.c_uint, vp] + [vp] * 14
    lib.moss_smpl_joint_tables.restype = C.c_int
    lib.moss_smpl_joint_tables.argtypes = [vp, C.c_int, C.c_int] + [vp] * 5
    lib.moss_smpl_frame_backward.restype = C.c_int
    lib.moss_smpl_frame_backward.argtypes = [vp, C.POINTER(MossSmplFrame)] + [vp] * 6
    lib.moss_image_losses_scratch_bytes.restype = C.c_size_t
    lib.moss_image_losses_scratch_bytes.argtypes = [C.c_int, C.c_int]
    lib.moss_sizeof_image_loss.restype = C.c_size_t
    lib.moss_image_losses.restype = C.c_int
    lib.moss_image_losses.argtypes = [vp, C.POINTER(MossImageLoss), vp, vp, vp]
    lib.moss_image_losses_static.restype = C.c_int
    lib.moss_image_losses_static.argtypes = [vp, C.POINTER(MossImageLoss), vp, vp, vp]
    lib.moss_mask_bounds.restype = C.c_int
    lib.moss_mask_bounds.argtypes = [vp, C.c_int, C.c_int, vp, vp]
    lib.moss_scale_gradients.restype = C.c_int
    lib.moss_scale_gradients.argtypes = [vp, vp, vp, C.c_size_t, vp, C.c_size_t]
    if lib.moss_sizeof_image_loss() != C.sizeof(MossImageLoss):
        raise RuntimeError("moss_b200: MossImageLoss layout mismatch between _abi.py and the library; rebuild")
    if lib.moss_sizeof_smpl_frame() != C.sizeof(MossSmplFrame):
        raise RuntimeError("moss_b200: MossSmplFrame layout mismatch between _abi.py and the library; rebuild")
    if lib.moss_sizeof_adam_group() != C.sizeof(MossAdamGroup):
        raise RuntimeError("MossAdamGroup layout differs between include/moss_splat.h and moss_b200/_abi.py")
    if lib.moss_sizeof_skin() != C.sizeof(MossSkin):
        raise RuntimeError("moss_b200: MossSkin layout mismatch between _abi.py and the library; rebuild")
    if lib.moss_sizeof_scene() != C.sizeof(MossScene):
        raise RuntimeError("moss_b200: MossScene layout mismatch between _abi.py and the library; rebuild")
    _lib = lib
    return lib


def profile_read():
    """-> {stage name: (total ms, launches)} accumulated since the last read (needs moss_profile_enable(1))."""
    lib = load()
    ms = (C.c_double * STAGE_COUNT)()
    n = (C.c_ulonglong * STAGE_COUNT)()
    check(lib.moss_profile_read(ms, n), "moss_profile_read")
    return {lib.moss_stage_name(i).decode(): (ms[i], int(n[i])) for i in range(STAGE_COUNT)}


_NO_GUARD = contextlib.nullcontext()


def stream_ptr(dev):
    """cudaStream_t (as an int) of torch's current stream on ``dev``.  ``torch.cuda.current_stream(dev).cuda_stream`` builds a
    Stream object for it (~9 us per call: 80 us of a 0.8 ms training iteration made of nine library calls)."""
    import torch
    idx = dev.index
    return torch._C._cuda_getCurrentRawStream(torch.cuda.current_device() if idx is None else idx)


def on_device(dev):
    """``with on_device(dev):`` makes ``dev`` the current CUDA device like ``torch.cuda.device(dev)``, at no cost when it
    already is (the context manager object costs ~5 us per use)."""
    import torch
    idx = dev.index
    if idx is None or idx == torch._C._cuda_getDevice():
        return _NO_GUARD
    return torch.cuda.device(dev)


def check(status, what="moss"):
    if status == MOSS_OK:
        return
    lib = load()
    msg = lib.moss_error_string(status).decode()
    if status == MOSS_ERR_CUDA:
        msg += f" (cudaError {lib.moss_last_cuda_error()})"
    raise RuntimeError(f"{what}: {msg}")
