"""CPU: the loss oracle (oracle/loss_oracle.py) reproduces what the REAL reference utils/loss_utils.py computed through
train_ZJU.py's expressions (tests/golden/loss_*.npz, made by tests/golden/make_loss_golden.py in the build container),
values and autograd gradients."""
import os

import numpy as np
import pytest
import torch

from oracle import loss_oracle as LO

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
LOSS_CASES = ["mid", "ragged", "small", "full"]


def load_loss_case(name):
    z = np.load(os.path.join(GOLD, f"loss_{name}.npz"))
    return {k: torch.from_numpy(np.asarray(z[k])) if getattr(z[k], "ndim", 0) > 0 else z[k].item() for k in z.files}


@pytest.mark.parametrize("name", LOSS_CASES)
def test_oracle_matches_reference(name):
    d = load_loss_case(name)
    image, alpha = d["image"].clone().requires_grad_(True), d["alpha"].clone().requires_grad_(True)
    assert LO.bounding_rect(d["bound"][0]) == tuple(d["rect"].tolist())
    total, t = LO.image_losses(image, alpha, d["gt"], d["bkgd"], d["bound"])
    for key, got in (("l1", t["Ll1"]), ("mask", t["mask_loss"]), ("ssim", t["ssim"]), ("s3im", t["s3im"]), ("total", total)):
        assert abs(float(got) - d[key]) <= 2e-6 * max(1.0, abs(d[key])), key
    gi, ga = torch.autograd.grad(total, [image, alpha])
    assert float((gi - d["d_total_dimage"]).abs().max()) <= 1e-6 * float(d["d_total_dimage"].abs().max())
    assert float((ga - d["d_total_dalpha"]).abs().max()) <= 1e-6 * float(d["d_total_dalpha"].abs().max())
