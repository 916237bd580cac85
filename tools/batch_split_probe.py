"""Developer probe: block shape (MOSS_FLAG_SPLIT_*) x segmented backward for a batched ZJU step.
    python tools/batch_split_probe.py [V]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import moss_b200 as M  # noqa: E402
from moss_b200 import _abi, synth  # noqa: E402
from tests import scenes as S  # noqa: E402

V = int(sys.argv[1]) if len(sys.argv) > 1 else 23
name = sys.argv[2] if len(sys.argv) > 2 else "zju"
dev = torch.device("cuda")
d = S.make_scene(name)
cam = d["cam"]
W, H = cam.image_width, cam.image_height
cams = synth.ring_cameras(23, W, H)[:V]
rs = [M.GaussianRasterizationSettings(**synth.settings_kwargs(c, d["bg"], d["sh_degree"], d["scale_modifier"], device=dev)) for c in cams]
to = lambda t: torch.empty(0) if t is None else t.to(dev)  # noqa: E731
inp = [to(d[k]) for k in ("means3D", "shs", "colors_precomp", "opacities", "scales", "rotations", "cov3D_precomp")]
gs = [S.upstream_grads(W, H, seed=7 + v, depth=False) for v in range(V)]
gc, gd, ga = (torch.stack([g[i] for g in gs]).to(dev) for i in range(3))
lib = _abi.load()
flush = torch.empty(64 * 1024 * 1024, dtype=torch.float32, device=dev)
for seg in (False, True):
    M.set_segmented_backward(seg)
    for split in (1, 2, 3):
        flags = split << _abi.FLAG_SPLIT_SHIFT
        for it in range(3):
            c, r, dp, a, st = M.forward_batch_raw(*inp, rs, flags=flags)
            M.backward_raw(st, r, gc, gd, ga)
        torch.cuda.synchronize()
        lib.moss_profile_enable(1)
        for it in range(5):
            flush.zero_()
            c, r, dp, a, st = M.forward_batch_raw(*inp, rs, flags=flags)
            M.backward_raw(st, r, gc, gd, ga)
        torch.cuda.synchronize()
        prof = _abi.profile_read()
        lib.moss_profile_enable(0)
        k = {n: round(v[0] / max(v[1], 1) * 1e3, 1) for n, v in prof.items() if v[1]}
        tot = sum(k.values())
        print(f"V={V} split={split} segmented={seg}: total {tot:.0f} us ({tot / V:.1f}/view)  " + " ".join(f"{n}={v}" for n, v in k.items()), flush=True)
