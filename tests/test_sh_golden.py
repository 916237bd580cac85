"""CPU: the oracle's SH evaluation (oracle/raster_oracle.py::sh_to_rgb, a restatement of DGR forward.cu:20-71) against
tests/golden/eval_sh.npz -- the reference's own python ``utils/sh_utils.py::eval_sh`` (:57-112) run by
tests/golden/make_sh_golden.py the way gaussian_renderer/__init__.py:100-105 calls it.  Same polynomial, different
association (python evaluates left to right without fma), so the tolerance is a few fp32 ulps of O(1) colours: 2e-6.
When /root/reference is mounted the live function is compared too."""
import importlib.util
import os

import numpy as np
import pytest
import torch

from oracle import raster_oracle as O
from tests import scenes as S

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "eval_sh.npz")
CASES = ["small", "small_d0", "small_d1", "small_d2", "ragged"]
SH_TOL = 2e-6


def _oracle_rgb(d):
    dirs = d["means3D"] - d["cam"].camera_center[None]
    return O.sh_to_rgb(d["sh_degree"], d["shs"], dirs)


@pytest.mark.parametrize("name", CASES)
def test_oracle_sh_matches_reference_eval_sh(name):
    z = np.load(GOLDEN)
    d = S.make_scene(name)
    raw = _oracle_rgb(d)
    assert float((raw - torch.from_numpy(z[name + "_raw"])).abs().max()) <= SH_TOL
    clamped = torch.clamp_min(raw, 0.0)
    assert float((clamped - torch.from_numpy(z[name])).abs().max()) <= SH_TOL


@pytest.mark.skipif(not os.path.exists("/root/reference/utils/sh_utils.py"), reason="reference tree not mounted")
def test_golden_file_is_what_the_live_reference_computes():
    spec = importlib.util.spec_from_file_location("ref_sh_utils", "/root/reference/utils/sh_utils.py")
    SH = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(SH)
    z = np.load(GOLDEN)
    d = S.make_scene("small")
    dirs = d["means3D"] - d["cam"].camera_center[None]
    live = SH.eval_sh(3, d["shs"].transpose(1, 2).reshape(-1, 3, 16), dirs / dirs.norm(dim=1, keepdim=True)) + 0.5
    assert np.array_equal(live.numpy().astype(np.float32), z["small_raw"])
