"""Generate tests/golden/*.npz: outputs of the UNMODIFIED reference rasterizer (oracle/_ref/libref_raster.so,
built by oracle/build_ref.py from /root/reference/submodules/diff-gaussian-rasterization) on the seeded CPU_SCENES.

Runs on the GPU box:  gpurun -- 'python tests/golden/make_golden.py gpurun_out/golden'   then copy the .npz here.
The reference has no CPU path and ships no golden vectors, so these files are the pin for
the CPU oracle (tests/test_oracle_golden.py) and for the CUDA product path (tests/test_parity_gpu.py).
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref_shim  # noqa: E402
from tests import scenes as S  # noqa: E402


def main(out_dir):
    os.makedirs(out_dir, exist_ok=True)
    for name in S.CPU_SCENES:
        d = S.make_scene(name)
        cam = d["cam"]
        ref = ref_shim.RefCall(d)
        color, radii, depth, alpha = ref.forward()
        vis = (radii > 0)
        rec = dict(radii=radii, tiles_touched=ref.state("tiles_touched"), keys_sorted=ref.state("keys_sorted"),
                   point_list=ref.state("point_list"), ranges=ref.state("ranges"), n_contrib=ref.state("n_contrib"),
                   final_T=ref.state("final_T"), color=color, depth=depth, alpha=alpha,
                   depths=ref.state("depths") * vis, means2D=ref.state("means2D") * vis[:, None],
                   conic_opacity=ref.state("conic_opacity") * vis[:, None],
                   num_rendered=torch.tensor(ref.R))
        if d["shs"] is not None:
            rec["rgb"] = ref.state("rgb") * vis[:, None]
            rec["clamped"] = ref.state("clamped") * vis[:, None]
        g = ref.backward(*S.upstream_grads(cam.image_width, cam.image_height))
        g2 = ref.backward(*S.upstream_grads(cam.image_width, cam.image_height))   # run-to-run atomic noise floor
        noise = {k: float((g[k] - g2[k]).double().norm() / max(float(g[k].double().norm()), 1e-30)) for k in g}
        for k, v in g.items():
            rec["g_" + k] = v
        rec = {k: v.detach().cpu().numpy() for k, v in rec.items()}
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), **rec)
        print(name, "R =", ref.R, "visible =", int(vis.sum()), "grad noise floor:",
              {k: f"{v:.1e}" for k, v in noise.items()}, flush=True)


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "golden"))
