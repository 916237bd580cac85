"""Generate tests/golden/eval_sh.npz by running the REAL reference ``utils/sh_utils.py::eval_sh`` (:57-112) on the CPU,
exactly as ``gaussian_renderer/__init__.py:100-105`` calls it when ``pipe.convert_SHs_python`` is set:

    shs_view = features.transpose(1, 2).view(-1, 3, (max_sh_degree + 1) ** 2)
    dir_pp_normalized = (means3D - camera_center) / |means3D - camera_center|
    colors_precomp = clamp_min(eval_sh(active_sh_degree, shs_view, dir_pp_normalized) + 0.5, 0)

Run in the build container only (needs /root/reference; the GPU box never reads it):

    python tests/golden/make_sh_golden.py

The file pins (a) the oracle's SH restatement (tests/test_sh_golden.py, CPU) and (b) the CUDA kernel's SH path
(tests/test_sh_gpu.py) to the reference's own python evaluation of the same polynomial.
"""
import importlib.util
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF = "/root/reference"
sys.path.insert(0, ROOT)
from tests import scenes as S  # noqa: E402

SH_CASES = {"small": 3, "small_d0": 0, "small_d1": 1, "small_d2": 2, "ragged": 3}


def main():
    spec = importlib.util.spec_from_file_location("ref_sh_utils", os.path.join(REF, "utils", "sh_utils.py"))
    SH = importlib.util.module_from_spec(spec)          # the real reference module, unmodified
    spec.loader.exec_module(SH)
    rec = {}
    for name, deg in SH_CASES.items():
        d = S.make_scene(name)
        assert d["sh_degree"] == deg
        shs_view = d["shs"].transpose(1, 2).reshape(-1, 3, 16)
        dir_pp = d["means3D"] - d["cam"].camera_center.repeat(d["shs"].shape[0], 1)
        dirn = dir_pp / dir_pp.norm(dim=1, keepdim=True)
        sh2rgb = SH.eval_sh(deg, shs_view, dirn)
        rec[name] = torch.clamp_min(sh2rgb + 0.5, 0.0).numpy().astype(np.float32)
        rec[name + "_raw"] = (sh2rgb + 0.5).numpy().astype(np.float32)
        print(name, "deg", deg, "clamped entries", int((sh2rgb + 0.5 < 0).sum()))
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "eval_sh.npz"), **rec)


if __name__ == "__main__":
    main()
