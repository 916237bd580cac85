"""Host-side mirror of the vendored ``diff_gaussian_rasterization`` Python API.

Same names, argument meaning, return tuple and error behaviour as
DGR/diff_gaussian_rasterization/__init__.py (``GaussianRasterizationSettings`` :160-172,
``GaussianRasterizer`` :174-223, ``rasterize_gaussians`` :21-42, ``_RasterizeGaussians`` :44-158), so
``gaussian_renderer/__init__.py::render()`` (:16, :36-55, :112-120) runs unchanged.  Underneath, torch only
provides device memory and the current stream; all compute is the sm_100a C-ABI library (no fallback).

Differences that are invisible through the API:
  * launches go to torch's CURRENT stream (the reference uses the legacy default stream);
  * no blocking D2H copy in the middle of the forward, no per-call zero-fills of outputs;
  * the three opaque state tensors are caller-owned workspaces (`geom+image` in one allocation, `binning`).
"""
from typing import NamedTuple
import contextlib
import ctypes as C
import os
import sys

import torch
import torch.nn as nn

from . import _abi


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


def _align(n, a=256):
    return (n + a - 1) // a * a


def _ptr(t):
    """Device pointer of a tensor, or NULL for the reference's empty-tensor placeholders."""
    if t is None or t.numel() == 0:
        return None
    return t.data_ptr()


def _f32c(t, name):
    if t.numel() == 0:
        return t
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor")
    if t.dtype != torch.float32:
        raise RuntimeError(f"{name} must be float32")
    return t.contiguous()


# remembered instance counts so that the binning workspace is right-sized without a host sync
_capacity_hint = {}
_EMPTY = torch.Tensor([])       # the reference's placeholder for an absent optional input (DGR/.../__init__.py:200-210)

# Segmented backward (checkpoint workspace, include/moss_splat.h MossScene.ckpt_ws): "auto" turns it on for images of
# up to 4096 tiles, where a few deep tiles set the backward's duration; larger images have enough tiles to fill the
# GPU and skip the extra checkpoint traffic.  True / False force it (tests compare both).
_segmented_backward = "auto"


def set_segmented_backward(mode):
    """mode: "auto" | True | False.  Returns the previous setting."""
    global _segmented_backward
    if mode not in ("auto", True, False):
        raise ValueError("mode must be 'auto', True or False")
    prev, _segmented_backward = _segmented_backward, mode
    return prev


def _want_ckpt(W, H, V=1):
    if _segmented_backward == "auto":
        T = ((W + 15) // 16) * ((H + 15) // 16)
        # a batch of many views already fills the GPU with tiles: the checkpoint traffic then costs more than the shorter
        # critical path saves (tools/batch_split_probe.py: 8 ZJU views 525 vs 539 us with segments, 23 views 1314 vs 1251)
        return T <= 4096 and T * V <= 12288
    return bool(_segmented_backward)


class RasterState:
    """Workspaces + scene descriptor of one forward call (kept alive for the backward)."""
    __slots__ = ("scene", "ws", "binning", "ckpt", "flags", "num_rendered", "keep", "grad_buf", "param_grad_numel",
                 "V", "views", "batched")

    def __init__(self):
        self.scene = _abi.MossScene()
        self.ws = None
        self.binning = None
        self.ckpt = None
        self.flags = 0
        self.num_rendered = 0
        self.keep = None
        self.grad_buf = None
        self.param_grad_numel = 0
        self.V = 1
        self.views = None          # ctypes array of MossView (batched calls)
        self.batched = False


# ---- asynchronous launch mode of the PUBLIC API (CUDA-graph capture of a whole training iteration, graphs.py) -----------
# Normally a forward waits for the instance count (like the reference, which copies num_rendered to the host in the middle
# of the call) so that the binning workspace can grow.  Inside `with async_launch():` the public entry points
# (GaussianRasterizer, rasterize_batch, fused.skin_and_rasterize) pass MOSS_FLAG_ASYNC instead: nothing waits, the binning
# capacity is 1.25 x the count the last synchronous call of the same shape produced (call once outside the block first),
# and an overflow only raises a flag on the device.  Every forward inside the block drops its int32[4] counters
# (num_rendered, overflow flag, capacity, non-empty tiles) into `async_counters()` for a later check.
_default_flags = 0
_async_counters = []


@contextlib.contextmanager
def async_launch():
    global _default_flags
    prev, _default_flags = _default_flags, _default_flags | _abi.FLAG_ASYNC
    try:
        yield
    finally:
        _default_flags = prev


def async_counters(clear=True):
    """The counter tensors of the forwards run under ``async_launch()`` since the last call (device tensors: reading them
    synchronises)."""
    global _async_counters
    out = _async_counters
    if clear:
        _async_counters = []
    return out


def _forward(means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, settings_list, flags, batched,
             skin=None):
    """``skin``: (filled _abi.MossSkin, tensors to keep alive, means3D_out or None, transforms_out or None) -- the fused
    skin + rasterize call (moss_skin_rasterize_forward); ``means3D`` then only names the device and the Gaussian count."""
    lib = _abi.load()
    rs = settings_list[0]
    V = len(settings_list)
    if means3D.ndim != 2 or means3D.shape[1] != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")   # DGR/rasterize_points.cu:57-59
    if not means3D.is_cuda:
        raise RuntimeError("moss_b200 rasterizer needs CUDA tensors (there is no CPU path)")
    if not 1 <= V <= _abi.MAX_VIEWS:
        raise RuntimeError(f"a batched call takes 1..{_abi.MAX_VIEWS} views, got {V}")
    dev = means3D.device
    means3D = _f32c(means3D, "means3D")
    sh = _f32c(sh, "sh"); colors_precomp = _f32c(colors_precomp, "colors_precomp")
    opacities = _f32c(opacities, "opacities"); scales = _f32c(scales, "scales")
    rotations = _f32c(rotations, "rotations"); cov3Ds_precomp = _f32c(cov3Ds_precomp, "cov3Ds_precomp")
    cams = []
    for r in settings_list:
        if (int(r.image_height), int(r.image_width), int(r.sh_degree), float(r.scale_modifier), bool(r.prefiltered)) != \
                (int(rs.image_height), int(rs.image_width), int(rs.sh_degree), float(rs.scale_modifier), bool(rs.prefiltered)):
            raise RuntimeError("all views of a batched call share image size, SH degree, scale modifier and prefiltered")
        cams.append((_f32c(r.bg, "bg"), _f32c(r.viewmatrix, "viewmatrix"), _f32c(r.projmatrix, "projmatrix"),
                     _f32c(r.campos, "campos")))

    P = means3D.shape[0]
    H, W = int(rs.image_height), int(rs.image_width)
    M = sh.shape[1] if sh.numel() != 0 else 0
    if any(r.debug for r in settings_list):
        flags |= _abi.FLAG_DEBUG
    export = 1 if (flags & _abi.FLAG_EXPORT_KEYS) else 0

    with _abi.on_device(dev):
        # three independent output tensors, like the reference (DGR/rasterize_points.cu:68-71): callers edit renders
        # in place (render_ZJU.py), which autograd refuses on views of a shared multi-output buffer
        lead = (V,) if batched else ()
        color = _alloc(lead + (3, H, W), torch.float32, dev)
        depth = _alloc(lead + (1, H, W), torch.float32, dev)
        alpha = _alloc(lead + (1, H, W), torch.float32, dev)
        radii = _alloc(lead + (P,), torch.int32, dev)
        gbytes = _align(lib.moss_geom_workspace_bytes_batch(P, V))
        ibytes = lib.moss_image_workspace_bytes_batch(W, H, V)
        ws = _alloc((gbytes + ibytes,), torch.uint8, dev)
        key = (P, W, H, V, dev.index)
        cap = _capacity_hint.get(key, 0)
        cap = max(int(cap * 1.25) + 4096, (4 * P + 4096) * V)
        st = RasterState()
        st.V, st.batched = V, batched
        sc = st.scene
        sc.P, sc.D, sc.M, sc.W, sc.H = P, int(rs.sh_degree), M, W, H
        sc.means3D = None if skin is not None else _ptr(means3D)
        sc.shs = _ptr(sh); sc.colors_precomp = _ptr(colors_precomp)
        sc.opacities = _ptr(opacities); sc.scales = _ptr(scales); sc.rotations = _ptr(rotations)
        sc.cov3D_precomp = _ptr(cov3Ds_precomp)
        sc.scale_modifier = float(rs.scale_modifier)
        sc.prefiltered = 1 if rs.prefiltered else 0
        if batched:
            views = (_abi.MossView * V)()
            for v, (r, (bg, view, proj, campos)) in enumerate(zip(settings_list, cams)):
                views[v].viewmatrix = _ptr(view); views[v].projmatrix = _ptr(proj); views[v].campos = _ptr(campos)
                views[v].bg = _ptr(bg); views[v].tanfovx = float(r.tanfovx); views[v].tanfovy = float(r.tanfovy)
            st.views = views
        else:
            bg, view, proj, campos = cams[0]
            sc.bg = _ptr(bg); sc.viewmatrix = _ptr(view); sc.projmatrix = _ptr(proj); sc.campos = _ptr(campos)
            sc.tanfovx = float(rs.tanfovx); sc.tanfovy = float(rs.tanfovy)
        sc.geom_ws = ws.data_ptr(); sc.geom_bytes = gbytes
        sc.image_ws = ws.data_ptr() + gbytes; sc.image_bytes = ibytes
        stream = _abi.stream_ptr(dev)
        nr = C.c_int(0)
        call_flags = flags
        want_ckpt = _want_ckpt(W, H, V)
        while True:
            bbytes = lib.moss_binning_workspace_bytes_ex(cap, export)
            binning = _alloc((bbytes,), torch.uint8, dev)
            sc.binning_ws = binning.data_ptr(); sc.binning_bytes = bbytes
            ckpt = None
            if want_ckpt:
                cbytes = lib.moss_ckpt_workspace_bytes_batch(lib.moss_binning_capacity(bbytes, export), W, H, V)
                ckpt = _alloc((cbytes,), torch.uint8, dev)
                sc.ckpt_ws = ckpt.data_ptr(); sc.ckpt_bytes = cbytes
            if skin is not None:
                status = lib.moss_skin_rasterize_forward(stream, C.byref(sc), C.byref(skin[0]), call_flags, color.data_ptr(),
                                                         depth.data_ptr(), alpha.data_ptr(), _ptr(radii), C.byref(nr),
                                                         _ptr(skin[2]) if skin[2] is not None else None,
                                                         _ptr(skin[3]) if skin[3] is not None else None)
            elif batched:
                status = lib.moss_rasterize_forward_batch(stream, C.byref(sc), V, st.views, call_flags, color.data_ptr(),
                                                          depth.data_ptr(), alpha.data_ptr(), _ptr(radii), C.byref(nr))
            else:
                status = lib.moss_rasterize_forward(stream, C.byref(sc), call_flags, color.data_ptr(), depth.data_ptr(),
                                                    alpha.data_ptr(), _ptr(radii), C.byref(nr))
            if status == _abi.MOSS_ERR_BINNING_OVERFLOW:
                cap = int(nr.value * 1.25) + 4096
                call_flags = flags | _abi.FLAG_REBIN
                continue
            _abi.check(status, "moss_rasterize_forward")
            break
        if nr.value >= 0:
            _capacity_hint[key] = nr.value
        st.ws, st.binning, st.ckpt, st.flags, st.num_rendered = ws, binning, ckpt, flags, nr.value
        # tensors whose device pointers sit in the scene struct must outlive the backward
        st.keep = (means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp) + tuple(t for c in cams for t in c)
        if skin is not None:
            st.keep = st.keep + (skin[0],) + tuple(skin[1])
    if (flags & _abi.FLAG_ASYNC) and (_default_flags & _abi.FLAG_ASYNC) and P > 0:
        _async_counters.append(export_state(st, "counters"))
    return color, radii, depth, alpha, st


def forward_raw(means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, raster_settings,
                flags=0):
    """Forward through the C ABI.  Returns (color, radii, depth, alpha, RasterState)."""
    return _forward(means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, [raster_settings], flags, False)


def forward_batch_raw(means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, settings_list, flags=0):
    """Batched forward (include/moss_splat.h: moss_rasterize_forward_batch): V cameras of the same Gaussians in one
    launch sequence.  Returns (color [V,3,H,W], radii [V,P], depth [V,1,H,W], alpha [V,1,H,W], RasterState)."""
    return _forward(means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, list(settings_list), flags, True)


def grad_layout(P, M, has_sh, has_sr, V=1):
    """Packed gradient buffer (floats): parameter gradients first -- the prefix a data-parallel job all-reduces
    -- then the blend-only accumulators.  With V > 1 views the screen-space gradient `means2D` is per view
    ([V,P,3], like `conic`) and leaves the all-reduced prefix.
    -> (dict name -> (offset, numel)), param_numel, zero_from, total."""
    lay, o = {}, 0
    pad = lambda n: (n + 3) // 4 * 4   # noqa: E731  every sub-array starts 16 B-aligned (vector stores of dsh / drot rows)
    for name, n in (("means3D", 3 * P), ("cov3D", 6 * P), ("sh", 3 * M * P if has_sh else 0),
                    ("scales", 3 * P if has_sr else 0), ("rotations", 4 * P if has_sr else 0)):
        lay[name] = (o, n); o += pad(n)
    zero_from = o
    lay["opacity"] = (o, P); o += pad(P)
    if V > 1:
        param_numel = o
    lay["means2D"] = (o, 3 * P * V); o += pad(3 * P * V)
    if V == 1:
        param_numel = o
    for name, n in (("colors", 3 * P), ("conic", 4 * P * V)):
        lay[name] = (o, n); o += pad(n)
    return lay, param_numel, zero_from, o


def _alloc(shape, dtype, dev):
    """Every buffer the kernels write is allocated here; tests/test_guard_gpu.py swaps in an allocator that surrounds each
    one with sentinel bytes (the stand-in for compute-sanitizer's memcheck, which is closed on this pool)."""
    return torch.empty(shape, dtype=dtype, device=dev)


_zeros = {}


def _shared_zeros(shape, dev):
    """A zero tensor shared by every caller (read-only by convention): the gradient of an input that was not given."""
    key = (tuple(shape), dev.index)
    t = _zeros.get(key)
    if t is None:
        if len(_zeros) > 64:
            _zeros.clear()
        t = _zeros[key] = torch.zeros(shape, dtype=torch.float32, device=dev)
    return t


def backward_raw(st, radii, grad_color, grad_depth, grad_alpha, flags=0, out=None, push=None):
    """Backward through the C ABI.  Returns the 8 gradient tensors in the reference's C++ order
    (DGR/rasterize_points.cu:205): dmeans2D, dcolors, dopacity, dmeans3D, dcov3D, dsh, dscales, drot.
    They are views of ONE packed buffer (kept in ``st.grad_buf``; ``st.param_grad_numel`` floats of parameter
    gradient lead it), which may be supplied as ``out``.  For a batched state the upstream gradients are
    [V,3,H,W] / [V,1,H,W], dmeans2D is [V,P,3] and the parameter gradients are summed over the views.
    ``push``: a ``parallel.PeerGradBuffer`` -- the data-parallel exchange is fused into the backward
    (include/moss_splat.h: moss_rasterize_backward_push): the gradients are written into its symmetric buffer and, once
    the queued work has run on every rank, the leading ``st.param_grad_numel`` floats hold the sum over the ranks."""
    lib = _abi.load()
    sc = st.scene
    P, M, H, W, V = sc.P, sc.M, sc.H, sc.W, st.V
    dev = radii.device
    lead = (V,) if st.batched else ()
    with _abi.on_device(dev):
        def _g(t, shape, required=False):
            if t is None:      # no gradient reaches this output: the ABI takes NULL for depth / alpha
                return torch.zeros(shape, dtype=torch.float32, device=dev) if required else None
            return _f32c(t, "grad").expand(shape).contiguous() if tuple(t.shape) != tuple(shape) else _f32c(t, "grad")
        grad_color = _g(grad_color, lead + (3, H, W), required=True)
        grad_depth = _g(grad_depth, lead + (1, H, W)); grad_alpha = _g(grad_alpha, lead + (1, H, W))
        has_sh = sc.shs is not None
        has_sr = sc.scales is not None
        lay, param_numel, zero_from, total = grad_layout(P, M, has_sh, has_sr, V)
        # every row of every output is written by the backward preprocess kernel: nothing is zero-filled here
        # (the reference fills 9 tensors, DGR/rasterize_points.cu:158-166); the atomically accumulated blend
        # gradients live in a self-cleaning scratch inside the geometry workspace
        if push is not None:
            if out is not None and out.data_ptr() != push.tensor.data_ptr():
                raise RuntimeError("with push= the gradients are written into the PeerGradBuffer's own tensor")
            if push.tensor.numel() < total:
                raise RuntimeError(f"PeerGradBuffer holds {push.tensor.numel()} floats, this call needs {total}")
            out = push.tensor
        buf = out if out is not None else _alloc((total,), torch.float32, dev)

        def v(name, *shape):
            o, n = lay[name]
            return buf[o:o + n].view(*shape)
        dmeans3D, dcov3D = v("means3D", P, 3), v("cov3D", P, 6)
        dopacity, dcolors = v("opacity", P, 1), v("colors", P, 3)
        dmeans2D, dconic = v("means2D", *lead, P, 3), v("conic", *lead, P, 2, 2)
        dsh = v("sh", P, M, 3) if has_sh else torch.zeros((P, 0, 3), dtype=torch.float32, device=dev)
        if has_sr:
            dscales, drot = v("scales", P, 3), v("rotations", P, 4)
        else:
            # absent inputs: the reference hands back zero tensors (DGR/rasterize_points.cu:158-166).  Two fill kernels per
            # backward (7 us of a 156 us ZJU step) for tensors nobody reads: shared, read-only zeros instead
            dscales, drot = _shared_zeros((P, 3), dev), _shared_zeros((P, 4), dev)
        stream = _abi.stream_ptr(dev)
        outs = (_ptr(dmeans2D), _ptr(dconic), _ptr(dopacity), _ptr(dcolors), _ptr(dmeans3D), _ptr(dcov3D),
                _ptr(dsh) if has_sh else None, _ptr(dscales) if has_sr else None, _ptr(drot) if has_sr else None)
        if push is not None:
            tbytes = lib.moss_batch_grad_workspace_bytes(P, M, V) if st.batched else 0
            tmp = _alloc((max(tbytes, 16),), torch.uint8, dev)
            status = lib.moss_rasterize_backward_push(
                stream, C.byref(sc), V, st.views if st.batched else None, flags | st.flags, _ptr(radii),
                grad_color.data_ptr(), _ptr(grad_depth), _ptr(grad_alpha), *outs, tmp.data_ptr(), tbytes,
                C.byref(push.descriptor(param_numel)))
        elif st.batched:
            tbytes = lib.moss_batch_grad_workspace_bytes(P, M, V)
            tmp = _alloc((max(tbytes, 16),), torch.uint8, dev)
            status = lib.moss_rasterize_backward_batch(
                stream, C.byref(sc), V, st.views, flags | st.flags, _ptr(radii),
                grad_color.data_ptr(), _ptr(grad_depth), _ptr(grad_alpha), *outs, tmp.data_ptr(), tbytes)
        else:
            status = lib.moss_rasterize_backward(
                stream, C.byref(sc), flags | st.flags, _ptr(radii),
                grad_color.data_ptr(), _ptr(grad_depth), _ptr(grad_alpha), *outs)
        _abi.check(status, "moss_rasterize_backward")
        st.grad_buf, st.param_grad_numel = buf, param_numel
    return dmeans2D, dcolors, dopacity, dmeans3D, dcov3D, dsh, dscales, drot


backward_batch_raw = backward_raw      # the state knows whether it came from a batched forward


# The public API runs through the compiled autograd binding (moss_b200/_C.so, csrc/torch_binding.cpp: the same glue as
# the Python autograd.Function below, written in C++ like the reference's DGR/rasterize_points.cu + ext.cpp) when it is
# built; MOSS_PUBLIC_API=python forces the Python Function, =cpp makes a missing binding an error.  Both drive the same
# C-ABI kernels -- neither is a fallback in the sense of a different compute path.
_binding = None
_binding_tried = False


def _load_binding():
    global _binding, _binding_tried
    if _binding_tried:
        return _binding
    _binding_tried = True
    mode = os.environ.get("MOSS_PUBLIC_API", "auto")
    if mode == "python":
        return None
    try:
        _abi.load()
        from . import _C as b
        if b.abi_version() != _abi.load().moss_abi_version():
            raise ImportError("moss_b200/_C.so was built against another libmoss_splat.so; rebuild (python -m moss_b200.build)")
        _binding = b
    except ImportError as e:
        if mode == "cpp":
            raise
        print(f"moss_b200: compiled autograd binding unavailable ({e}); using the Python autograd.Function", file=sys.stderr)
    return _binding


def public_api_kind():
    return "compiled autograd binding (moss_b200/_C.so)" if _load_binding() is not None else "python autograd.Function over ctypes"


def _rasterize(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, raster_settings, batched):
    b = _load_binding() if not (_default_flags & _abi.FLAG_ASYNC) else None
    if b is None:
        return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                         cov3Ds_precomp, list(raster_settings) if batched else raster_settings)
    lst = list(raster_settings) if batched else [raster_settings]
    rs = lst[0]
    H, W = int(rs.image_height), int(rs.image_width)
    if batched:
        for r in lst[1:]:
            if (int(r.image_height), int(r.image_width), int(r.sh_degree), float(r.scale_modifier), bool(r.prefiltered)) != \
                    (H, W, int(rs.sh_degree), float(rs.scale_modifier), bool(rs.prefiltered)):
                raise RuntimeError("all views of a batched call share image size, SH degree, scale modifier and prefiltered")
    out = b.rasterize(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                      [r.bg for r in lst], [r.viewmatrix for r in lst], [r.projmatrix for r in lst], [r.campos for r in lst],
                      [float(r.tanfovx) for r in lst], [float(r.tanfovy) for r in lst], H, W, float(rs.scale_modifier),
                      int(rs.sh_degree), bool(rs.prefiltered), any(r.debug for r in lst), batched, _want_ckpt(W, H, len(lst)), 0)
    return tuple(out)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings):
    return _rasterize(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, raster_settings, False)


# Data-parallel training through the public API: with a sink set (parallel.PeerGradBuffer), the autograd backward writes
# its gradients into the sink's symmetric buffer with the exchange fused in (one exchange per backward, the gradient
# tensors handed to autograd are views of that buffer and hold the SUM over the ranks once the stream has run).
_gradient_sink = None


def set_gradient_sink(peer):
    """peer: parallel.PeerGradBuffer or None.  Returns the previous sink."""
    global _gradient_sink
    prev, _gradient_sink = _gradient_sink, peer
    b = _load_binding()
    if b is not None:
        if peer is None:
            b.set_gradient_sink(None, 0, 1, [], [], [], 0, 0, 0)
        else:
            g, st, pads = peer.raw_pointers()
            b.set_gradient_sink(peer.tensor, peer.rank, peer.world, g, st, pads, int(peer.handle.signal_pad_size),
                                int(peer.multicast_ptr or 0), peer.error.data_ptr())
    return prev


class _RasterizeGaussians(torch.autograd.Function):
    """One view (``raster_settings`` is a GaussianRasterizationSettings) or a batch (a list/tuple of them)."""

    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                raster_settings):
        batched = isinstance(raster_settings, (list, tuple)) and not isinstance(raster_settings, GaussianRasterizationSettings)
        color, radii, depth, alpha, st = _forward(means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                                                  list(raster_settings) if batched else [raster_settings], _default_flags, batched)
        ctx.raster_settings = raster_settings
        ctx.num_rendered = st.num_rendered
        ctx.mark_non_differentiable(radii)
        ctx.set_materialize_grads(False)       # an output nobody differentiates arrives as None, not as a zero image
        # The workspaces and the inputs whose device pointers sit in st.scene go through save_for_backward, like
        # the reference's three state tensors (DGR/diff_gaussian_rasterization/__init__.py:97): autograd frees
        # them right after the backward unless retain_graph is set.
        ctx.save_for_backward(radii, st.ws, st.binning, *st.keep, *([st.ckpt] if st.ckpt is not None else []))
        st.ws = st.binning = st.keep = st.ckpt = None
        ctx.state = st
        return color, radii, depth, alpha

    @staticmethod
    def backward(ctx, grad_out_color, grad_radii, grad_depth, grad_alpha):
        saved = ctx.saved_tensors           # raises the usual error when the graph was already freed
        radii = saved[0]
        st = ctx.state
        (grad_means2D, grad_colors_precomp, grad_opacities, grad_means3D, grad_cov3Ds_precomp, grad_sh,
         grad_scales, grad_rotations) = backward_raw(st, radii, grad_out_color, grad_depth, grad_alpha, push=_gradient_sink)
        # same order as DGR/diff_gaussian_rasterization/__init__.py:146-156
        return (grad_means3D, grad_means2D, grad_sh, grad_colors_precomp, grad_opacities, grad_scales,
                grad_rotations, grad_cov3Ds_precomp, None)


def rasterize_batch(means3D, means2D, opacities, settings_list, shs=None, colors_precomp=None, scales=None,
                    rotations=None, cov3D_precomp=None):
    """V cameras of the SAME Gaussians in one launch sequence (no reference counterpart: MOSS renders one view per
    iteration, train_ZJU.py:93-100).  ``settings_list``: V GaussianRasterizationSettings that
    share image size, SH degree and scale modifier; ``means2D`` [V,P,3] is the screen-space gradient sink of every view.
    -> (color [V,3,H,W], radii [V,P], depth [V,1,H,W], alpha [V,1,H,W]); view v equals the single-view call bit for
    bit, and the parameter gradients are the sums over the views (in view order)."""
    if (shs is None) == (colors_precomp is None):
        raise Exception('Please provide excatly one of either SHs or precomputed colors!')
    if ((scales is None or rotations is None) and cov3D_precomp is None) or \
            ((scales is not None or rotations is not None) and cov3D_precomp is not None):
        raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
    e = _EMPTY
    return _rasterize(means3D, means2D, e if shs is None else shs, e if colors_precomp is None else colors_precomp,
                      opacities, e if scales is None else scales, e if rotations is None else rotations,
                      e if cov3D_precomp is None else cov3D_precomp, list(settings_list), True)


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        """bool[P]: view-space z > 0.2 (DGR/diff_gaussian_rasterization/__init__.py:179-188)."""
        lib = _abi.load()
        with torch.no_grad():
            rs = self.raster_settings
            pos = _f32c(positions, "positions")
            if not pos.is_cuda:
                raise RuntimeError("moss_b200 rasterizer needs CUDA tensors (there is no CPU path)")
            P = pos.shape[0]
            present = torch.empty((P,), dtype=torch.bool, device=pos.device)
            view = _f32c(rs.viewmatrix, "viewmatrix"); proj = _f32c(rs.projmatrix, "projmatrix")
            with _abi.on_device(pos.device):
                status = lib.moss_mark_visible(_abi.stream_ptr(pos.device), P, _ptr(pos),
                                               _ptr(view), _ptr(proj), _ptr(present))
            _abi.check(status, "moss_mark_visible")
        return present

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None):
        raster_settings = self.raster_settings
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        empty = _EMPTY
        shs = empty if shs is None else shs
        colors_precomp = empty if colors_precomp is None else colors_precomp
        scales = empty if scales is None else scales
        rotations = empty if rotations is None else rotations
        cov3D_precomp = empty if cov3D_precomp is None else cov3D_precomp
        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales, rotations,
                                   cov3D_precomp, raster_settings)


def export_state(st, name, view=None):
    """Copy one internal state array of a forward call into a fresh tensor (parity tests).  For a batched state
    ``view=v`` returns that view's slice in single-view conventions (ranges / point_list relative to the view's own
    instance list, Gaussian ids 0..P-1), ``view=None`` the raw batch arrays."""
    lib = _abi.load()
    sc = st.scene
    V = st.V
    P, N = sc.P * V, sc.W * sc.H * V
    T1 = ((sc.W + 15) // 16) * ((sc.H + 15) // 16)
    T = T1 * V
    R = max(st.num_rendered, 0)
    dev = st.ws.device
    spec = {
        "depths": (torch.float32, (P,)), "clamped": (torch.uint8, (P, 3)), "means2D": (torch.float32, (P, 2)),
        "cov3D": (torch.float32, (P, 6)), "conic_opacity": (torch.float32, (P, 4)), "rgb": (torch.float32, (P, 3)),
        "tiles_touched": (torch.int32, (P,)), "keys_sorted": (torch.int64, (R,)), "point_list": (torch.int32, (R,)),
        "ranges": (torch.int32, (T, 2)), "n_contrib": (torch.int32, (N,)), "final_T": (torch.float32, (N,)),
        "counters": (torch.int32, (4,)), "view_base": (torch.int32, (V + 1,)),
    }[name]
    out = torch.empty(spec[1], dtype=spec[0], device=dev)
    if out.numel() == 0:                 # nothing rendered (or P == 0): an empty list, no call with a null destination
        return out
    with _abi.on_device(dev):
        n = lib.moss_export_state_batch(_abi.stream_ptr(dev), C.byref(sc), V, _abi.STATE[name],
                                        out.data_ptr(), out.numel() * out.element_size())
    if n < 0:
        raise RuntimeError(f"moss_export_state({name}) failed: {lib.moss_error_string(int(-n)).decode()}")
    if view is None or name in ("counters", "view_base"):
        return out
    v = int(view)
    if name in ("keys_sorted", "point_list", "ranges"):
        base = export_state(st, "view_base").cpu()
        lo, hi = int(base[v]), int(base[v + 1])
        if name == "ranges":
            r = out[v * T1:(v + 1) * T1].clone()
            r[r[:, 1] > 0] -= lo                   # empty tiles stay (0,0) like the reference's memset
            return r
        if name == "point_list":
            return out[lo:hi] - v * sc.P
        return out[lo:hi].clone()
    per = out.shape[0] // V
    return out[v * per:(v + 1) * per].clone()


def export_state_counters(st):
    """int32[4] on the host side of a sync: num_rendered, overflow flag, capacity, non-empty tiles."""
    return export_state(st, "counters").cpu()


def add_densification_stats(radii, viewspace_grad, xyz_gradient_accum=None, denom=None, max_radii2D=None):
    """In place, for the Gaussians with ``radii > 0`` (the reference's visibility_filter): ``max_radii2D = max(max_radii2D,
    radii)`` (train_ZJU.py:173), ``xyz_gradient_accum += |viewspace_grad[:, :2]|`` and ``denom += 1``
    (GaussianModel.add_densification_stats, scene/gaussian_model.py:816-818) -- one kernel, no boolean-mask indexing
    (each of the reference's three masked updates synchronises the host through ``nonzero``).
    ``viewspace_grad`` is ``viewspace_point_tensor.grad`` ([P,3] float32); the statistics are float32 [P] or [P,1]."""
    lib = _abi.load()
    if not radii.is_cuda:
        raise RuntimeError("moss_b200 needs CUDA tensors (there is no CPU path)")
    P = radii.shape[0]
    if radii.dtype != torch.int32 or not radii.is_contiguous():
        raise RuntimeError("radii must be a contiguous int32 tensor (as returned by the rasterizer)")

    def stat(t, name):
        if t is None:
            return None
        if t.dtype != torch.float32 or not t.is_contiguous() or t.numel() != P or t.device != radii.device:
            raise RuntimeError(f"{name} must be a contiguous float32 tensor with one entry per Gaussian on the device of radii")
        return t
    acc, den, mx = stat(xyz_gradient_accum, "xyz_gradient_accum"), stat(denom, "denom"), stat(max_radii2D, "max_radii2D")
    g = None
    if acc is not None:
        g = _f32c(viewspace_grad, "viewspace_grad")
        if g.shape != (P, 3) or g.device != radii.device:
            raise RuntimeError("viewspace_grad must be [P,3] on the device of radii")
    with _abi.on_device(radii.device):
        status = lib.moss_densify_stats(_abi.stream_ptr(radii.device), P, _ptr(radii), _ptr(g),
                                        _ptr(acc), _ptr(den), _ptr(mx))
    _abi.check(status, "moss_densify_stats")
