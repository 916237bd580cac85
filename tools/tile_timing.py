"""Developer tool: per-tile phase timings (clock64) of the blend kernels from a -DMOSS_TILE_TIMING build."""
import os, sys, subprocess, ctypes as C
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from moss_b200 import build as B, _abi
dbg = os.path.join(ROOT, "moss_b200", "_lib", "libmoss_splat_dbg.so")
if "--build" in sys.argv:
    subprocess.check_call(["nvcc", *B.NVCC_FLAGS, "-DMOSS_TILE_TIMING", os.path.join(B.CSRC, "moss_splat.cu"), "-o", dbg])
    sys.exit(0)
_abi.LIB_PATH = dbg
import moss_b200 as M
from moss_b200 import synth
from tests import scenes as S
lib = _abi.load()
name = sys.argv[1] if len(sys.argv) > 1 else "zju"
dev = torch.device("cuda", 0)
d = S.make_scene(name)
cam = d["cam"]; W, H = cam.image_width, cam.image_height
to = lambda t: torch.empty(0) if t is None else t.to(dev)
rs = M.GaussianRasterizationSettings(**synth.settings_kwargs(cam, d["bg"], d["sh_degree"], d["scale_modifier"], device=dev))
gr = [t.to(dev) for t in S.upstream_grads(W, H, depth=False)]
T = ((W + 15) // 16) * ((H + 15) // 16)
split = int(sys.argv[2]) if len(sys.argv) > 2 else 0
flags = split << _abi.FLAG_SPLIT_SHIFT
ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
for it in range(4):
    ev[0].record()
    color, radii, depth, alpha, st = M.forward_raw(to(d["means3D"]), to(d["shs"]), to(d["colors_precomp"]), to(d["opacities"]),
                                                   to(d["scales"]), to(d["rotations"]), to(d["cov3D_precomp"]), rs, flags=flags)
    ev[1].record()
    M.backward_raw(st, radii, *gr)
    ev[2].record()
torch.cuda.synchronize()
print(f"{name} split={split}: forward {ev[0].elapsed_time(ev[1])*1e3:.1f} us, backward {ev[1].elapsed_time(ev[2])*1e3:.1f} us (events, warm L2, incl. host gaps)")
buf = np.zeros((T, 8), dtype=np.int64)
lib.moss_debug_tile_timing.argtypes = [C.c_void_p, C.c_int]
assert lib.moss_debug_tile_timing(buf.ctypes.data, T) == 0
det = np.zeros((T, 16), dtype=np.int64)
lib.moss_debug_tile_detail.argtypes = [C.c_void_p, C.c_int]
assert lib.moss_debug_tile_detail(det.ctypes.data, T) == 0
n = buf[:, 7]; last = buf[:, 6]
sort_ns, blend_ns, bwd_ns = buf[:, 1] - buf[:, 0], buf[:, 3] - buf[:, 2], buf[:, 5] - buf[:, 4]
busy = n > 0
print("times in ns (global timer); blend/bwd are sub-block 0 of each tile")
print(" tile     n  last   sort_ns  blend_ns    bwd_ns | sort steps (cycles): splitters search scatter rank")
for t in np.argsort(-n)[:12]:
    print(f"{t:5d} {n[t]:5d} {last[t]:5d} {sort_ns[t]:9d} {blend_ns[t]:9d} {bwd_ns[t]:9d} | {det[t,4]-det[t,8]:6d} {det[t,5]-det[t,4]:6d} {det[t,6]-det[t,5]:6d} {det[t,9]-det[t,6]:6d}")
print("launch spans (first start to last end, ns): sort", buf[busy, 1].max() - buf[busy, 0].min(), " blend fwd", buf[busy, 3].max() - buf[busy, 2].min(),
      " blend bwd", buf[busy, 5].max() - buf[busy, 4].min())
print("slowest: sort", np.sort(sort_ns[busy])[-3:], "blend", np.sort(blend_ns[busy])[-3:], "bwd", np.sort(bwd_ns[busy])[-3:])
