"""TEST/BENCH INFRASTRUCTURE ONLY -- ctypes driver of oracle/_ref/libref_raster.so, the UNMODIFIED reference
rasterizer (DGR/cuda_rasterizer/*.cu) behind our C-ABI shim (oracle/ref_shim.cu).  Needs a GPU.

Never imported by the product package ``moss_b200``.  Used by tests/ (bit-exact parity on the GPU box),
tests/golden/make_golden.py (golden-vector generation) and bench.py's reference-CUDA timing leg.
The reference launches on the legacy default stream and blocks on a D2H copy; callers synchronise around it.
"""
import ctypes as C
import os

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "_ref", "libref_raster.so")

_lib = None

STATE = dict(depths=(0, torch.float32, 1), clamped=(1, torch.uint8, 3), means2D=(2, torch.float32, 2),
             cov3D=(3, torch.float32, 6), conic_opacity=(4, torch.float32, 4), rgb=(5, torch.float32, 3),
             point_offsets=(6, torch.int32, 1), tiles_touched=(7, torch.int32, 1))
BIN_STATE = dict(keys_unsorted=(8, torch.int64), keys_sorted=(9, torch.int64), list_unsorted=(10, torch.int32),
                 point_list=(11, torch.int32))


def available():
    return os.path.exists(LIB_PATH)


def load():
    global _lib
    if _lib is None:
        lib = C.CDLL(LIB_PATH)
        vp, f, i = C.c_void_p, C.c_float, C.c_int
        lib.ref_forward.restype = i
        lib.ref_forward.argtypes = [i, i, i, vp, i, i, vp, vp, vp, vp, vp, f, vp, vp, vp, vp, vp, f, f, i,
                                    vp, vp, vp, vp, i]
        lib.ref_backward.restype = None
        lib.ref_backward.argtypes = [i, i, i, i, vp, i, i, vp, vp, vp, vp, vp, f, vp, vp, vp, vp, vp, f, f, vp,
                                     vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i]
        lib.ref_mark_visible.restype = None
        lib.ref_mark_visible.argtypes = [i, vp, vp, vp, vp]
        lib.ref_export.restype = C.c_longlong
        lib.ref_export.argtypes = [i, vp, C.c_longlong]
        lib.ref_last_cuda_error.restype = i
        _lib = lib
    return _lib


def _p(t):
    return None if t is None else t.data_ptr()


class RefCall:
    """One forward (+ optional backward) of the reference on CUDA tensors."""

    def __init__(self, d, device="cuda"):
        """d: scene dict of tests/scenes.py (CPU or CUDA tensors)."""
        self.lib = load()
        dev = torch.device(device)
        to = lambda t: None if t is None else t.to(dev, torch.float32).contiguous()  # noqa: E731
        cam = d["cam"]
        self.means3D, self.shs, self.colors = to(d["means3D"]), to(d["shs"]), to(d["colors_precomp"])
        self.opac, self.scales, self.rots, self.cov = to(d["opacities"]), to(d["scales"]), to(d["rotations"]), to(d["cov3D_precomp"])
        self.view, self.proj, self.campos = to(cam.world_view_transform), to(cam.full_proj_transform), to(cam.camera_center)
        self.bg = to(d["bg"])
        self.W, self.H = int(cam.image_width), int(cam.image_height)
        self.tanfovx, self.tanfovy = float(cam.tanfovx), float(cam.tanfovy)
        self.P = self.means3D.shape[0]
        self.D = int(d["sh_degree"])
        self.M = self.shs.shape[1] if self.shs is not None else 0
        self.mod = float(d["scale_modifier"])
        self.dev = dev
        self.R = 0

    def forward(self):
        dev, H, W, P = self.dev, self.H, self.W, self.P
        self.color = torch.zeros((3, H, W), device=dev)
        self.depth = torch.zeros((1, H, W), device=dev)
        self.alpha = torch.zeros((1, H, W), device=dev)
        self.radii = torch.zeros((P,), dtype=torch.int32, device=dev)
        torch.cuda.synchronize()
        self.R = self.lib.ref_forward(P, self.D, self.M, _p(self.bg), W, H, _p(self.means3D), _p(self.shs),
                                      _p(self.colors), _p(self.opac), _p(self.scales), self.mod, _p(self.rots),
                                      _p(self.cov), _p(self.view), _p(self.proj), _p(self.campos), self.tanfovx,
                                      self.tanfovy, 0, _p(self.color), _p(self.depth), _p(self.alpha),
                                      _p(self.radii), 0)
        torch.cuda.synchronize()
        return self.color, self.radii, self.depth, self.alpha

    def backward(self, gc, gd, ga):
        dev, P, M = self.dev, self.P, self.M
        z = lambda *s: torch.zeros(s, device=dev)  # noqa: E731
        gc, gd, ga = (t.to(dev, torch.float32).contiguous() for t in (gc, gd, ga))
        # The reference reads dL_ddepths[pix_id] / dL_dalphas[pix_id] for threads OUTSIDE the image too (missing braces at
        # DGR/cuda_rasterizer/backward.cu:460-464): up to 15 rows past the end when H or W is not a multiple of 16.  With a
        # tensor at the end of an allocator segment that is an illegal address (tests/test_zy_fuzz_gpu.py met it), so the two
        # arrays are handed over with that many floats of slack behind them.  The values read there are never used.
        slack = self.W * (((self.H + 15) // 16) * 16) + 16
        def padded(t):
            buf = torch.zeros(slack, device=dev)
            buf[: t.numel()] = t.reshape(-1)
            return buf
        gd, ga = padded(gd), padded(ga)
        g = dict(means2D=z(P, 3), conic=z(P, 2, 2), opacity=z(P, 1), colors=z(P, 3), means3D=z(P, 3), cov3D=z(P, 6),
                 sh=z(P, M, 3), scales=z(P, 3), rotations=z(P, 4))
        torch.cuda.synchronize()
        self.lib.ref_backward(P, self.D, M, self.R, _p(self.bg), self.W, self.H, _p(self.means3D), _p(self.shs),
                              _p(self.colors), _p(self.alpha), _p(self.scales), self.mod, _p(self.rots), _p(self.cov),
                              _p(self.view), _p(self.proj), _p(self.campos), self.tanfovx, self.tanfovy,
                              _p(self.radii), _p(gc), _p(gd), _p(ga), _p(g["means2D"]), _p(g["conic"]),
                              _p(g["opacity"]), _p(g["colors"]), _p(g["means3D"]), _p(g["cov3D"]), _p(g["sh"]),
                              _p(g["scales"]), _p(g["rotations"]), 0)
        torch.cuda.synchronize()
        return g

    def state(self, name):
        """Internal state of the last forward (the reference's GeometryState/BinningState/ImageState arrays)."""
        P, R, N = self.P, self.R, self.W * self.H
        T = ((self.W + 15) // 16) * ((self.H + 15) // 16)
        if name in STATE:
            which, dt, k = STATE[name]
            out = torch.empty((P, k) if k > 1 else (P,), dtype=dt, device=self.dev)
        elif name in BIN_STATE:
            which, dt = BIN_STATE[name]
            out = torch.empty((R,), dtype=dt, device=self.dev)
        elif name == "ranges":
            which, out = 12, torch.empty((T, 2), dtype=torch.int32, device=self.dev)
        elif name == "n_contrib":
            which, out = 13, torch.empty((N,), dtype=torch.int32, device=self.dev)
        elif name == "final_T":
            which, out = 14, torch.empty((N,), dtype=torch.float32, device=self.dev)
        else:
            raise KeyError(name)
        nbytes = out.numel() * out.element_size()
        if nbytes == 0:
            return out
        n = self.lib.ref_export(which, out.data_ptr(), nbytes)
        if n != nbytes:
            raise RuntimeError(f"ref_export({name}) returned {n}, expected {nbytes}")
        torch.cuda.synchronize()
        return out
