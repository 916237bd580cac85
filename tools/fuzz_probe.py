"""Developer tool: one scene spec (tests/scenes.py options, JSON on the command line) through the reference and this
library, everything printed side by side."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import moss_b200 as M
from tests import scenes as S
from tests.test_parity_gpu import run_ours
from oracle import ref_shim
spec = json.loads(sys.argv[1])
for trial in range(int(sys.argv[2]) if len(sys.argv) > 2 else 2):
    d = S.make_scene_from_spec(spec)
    cam = d["cam"]
    grads = S.upstream_grads(cam.image_width, cam.image_height, seed=11 + spec["P"], depth=True)
    ref = ref_shim.RefCall(d)
    rc, rr, rd, ra = ref.forward()
    o = run_ours(d, grads=grads)
    st = o["st"]
    print("trial", trial, "R ours/ref", o["R"], ref.R, "radii", o["radii"].tolist()[:8], rr.tolist()[:8])
    for nm in ("tiles_touched", "depths", "means2D", "conic_opacity", "keys_sorted", "point_list", "ranges"):
        a, b = M.export_state(st, nm), ref.state(nm)
        print(" ", nm, "equal" if torch.equal(a, b) else "DIFF", a.flatten().tolist()[:8], b.flatten().tolist()[:8])
    print("  color max diff", float((o["color"] - rc).abs().max()), "alpha sum", float(o["alpha"].sum()), float(ra.sum()))
    rg = ref.backward(*grads)
    for nm in ("means2D", "opacity", "means3D", "cov3D", "sh", "scales", "rotations"):
        if nm in o["grads"] and o["grads"][nm] is not None and o["grads"][nm].numel():
            print("  grad", nm, o["grads"][nm].flatten().tolist()[:6], rg[nm].flatten().tolist()[:6])
