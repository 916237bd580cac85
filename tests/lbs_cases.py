"""Shared loader for the LBS golden cases (tests/golden/lbs_*.npz, generated from the real reference)."""
import os

import numpy as np
import torch

from moss_b200 import synth

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
LBS_CASES = ["plain", "delta", "correct"]


def oracle_smpl(body):
    from oracle import lbs_oracle as L
    V = body.verts.shape[0]
    return L.Smpl(body.verts, body.shapedirs, body.posedirs.t().reshape(V, 3, 207).contiguous(), body.J_regressor,
                  body.weights, body.parents)


def load_case(name):
    """-> (dict of torch tensors from the npz, FrameTables built by the oracle, Body)."""
    from oracle import lbs_oracle as L
    z = np.load(os.path.join(GOLD, f"lbs_{name}.npz"))
    d = {k: torch.from_numpy(np.asarray(z[k])) if getattr(z[k], "ndim", 0) > 0 else z[k].item() for k in z.files}
    body = synth.body()
    fr = synth.frame_params(int(d["seed"]))
    cr = d["correct_Rs"].reshape(1, 23, 3, 3) if "correct_Rs" in d else None
    ft = L.frame_tables(oracle_smpl(body), fr.params, fr.big_params, cr)
    return d, ft, body
