"""GPU: no kernel writes outside the buffers it was handed.  compute-sanitizer is closed on this
pool, so every buffer of a forward + backward -- outputs, geometry / image / binning / checkpoint workspaces at exactly
the sizes the moss_*_workspace_bytes queries return, the packed gradient, the batch scratch -- is carved out of a larger
allocation filled with sentinel bytes; after the call the 4 KB in front of and behind each buffer must be untouched.
Scenes: ragged image and Gaussian counts, behind-camera culls, the Stress layout (whole-tile blocks, no checkpoints), a
3-view batch, both accumulation modes, and a forced binning overflow + regrow."""
import pytest
import torch

import moss_b200 as M
from moss_b200 import _abi, rasterizer as R, synth
from tests import scenes as S

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda:0")
GUARD = 4096
SENTINEL = 0xA5


class Guarded:
    def __init__(self):
        self.bufs = []

    def __call__(self, shape, dtype, dev):
        n = 1
        for d in shape:
            n *= int(d)
        nbytes = n * torch.empty((), dtype=dtype).element_size()
        flat = torch.full((nbytes + 2 * GUARD,), SENTINEL, dtype=torch.uint8, device=dev)
        self.bufs.append((flat, nbytes))
        return flat[GUARD:GUARD + nbytes].view(dtype).view(tuple(int(d) for d in shape))

    def check(self):
        torch.cuda.synchronize()
        assert self.bufs
        for i, (flat, nbytes) in enumerate(self.bufs):
            front, back = flat[:GUARD], flat[GUARD + nbytes:]
            assert bool((front == SENTINEL).all()), f"buffer {i} ({nbytes} B): bytes in FRONT of it were written"
            assert bool((back == SENTINEL).all()), f"buffer {i} ({nbytes} B): bytes BEHIND it were written"
        return len(self.bufs)


@pytest.fixture
def guarded(monkeypatch):
    g = Guarded()
    monkeypatch.setattr(R, "_alloc", g)
    return g


def _tensors(d):
    return {k: (d[k].to(DEV) if d[k] is not None else torch.empty(0)) for k in
            ("means3D", "shs", "colors_precomp", "opacities", "scales", "rotations", "cov3D_precomp")}


@pytest.mark.parametrize("name", ["ragged", "small_cov", "small_color", "big_splats", "monocap_edge", "stress_small"])
@pytest.mark.parametrize("det", [False, True])
def test_single_view_stays_inside_its_buffers(name, det, guarded):
    d = S.make_scene(name)
    cam = d["cam"]
    rs = M.GaussianRasterizationSettings(**synth.settings_kwargs(cam, d["bg"], d["sh_degree"], d["scale_modifier"], device=DEV))
    t = _tensors(d)
    gc, gd, ga = (x.to(DEV) for x in S.upstream_grads(cam.image_width, cam.image_height, depth=True))
    c, r, dp, a, st = M.forward_raw(t["means3D"], t["shs"], t["colors_precomp"], t["opacities"], t["scales"], t["rotations"],
                                    t["cov3D_precomp"], rs)
    M.backward_raw(st, r, gc, gd, ga, flags=_abi.FLAG_DETERMINISTIC if det else 0)
    assert guarded.check() >= 7
    assert int((r > 0).sum()) > 0


def test_batch_and_regrow_stay_inside_their_buffers(guarded):
    d = S.make_scene("ragged")
    W, H = d["cam"].image_width, d["cam"].image_height
    cams = synth.ring_cameras(23, W, H)[3:6]
    rs = [M.GaussianRasterizationSettings(**synth.settings_kwargs(c, d["bg"], d["sh_degree"], d["scale_modifier"], device=DEV)) for c in cams]
    t = _tensors(d)
    gs = [S.upstream_grads(W, H, seed=7 + v, depth=True) for v in range(3)]
    gc, gd, ga = (torch.stack([g[i] for g in gs]).to(DEV) for i in range(3))
    R._capacity_hint.clear()
    c, r, dp, a, st = M.forward_batch_raw(t["means3D"], t["shs"], t["colors_precomp"], t["opacities"], t["scales"], t["rotations"],
                                          t["cov3D_precomp"], rs)
    M.backward_batch_raw(st, r, gc, gd, ga)
    n1 = guarded.check()
    # first call of a scene whose splats cover dozens of tiles each: the initial capacity (4 P + 4096) overflows, the
    # forward learns R and re-runs the binning with a larger buffer
    R._capacity_hint.clear()
    big = S.make_scene("big_splats")
    tb = _tensors(big)
    rsb = M.GaussianRasterizationSettings(**synth.settings_kwargs(big["cam"], big["bg"], big["sh_degree"], big["scale_modifier"], device=DEV))
    c, r, dp, a, st = M.forward_raw(tb["means3D"], tb["shs"], tb["colors_precomp"], tb["opacities"], tb["scales"], tb["rotations"],
                                    tb["cov3D_precomp"], rsb)
    g3 = [x.to(DEV) for x in S.upstream_grads(big["cam"].image_width, big["cam"].image_height, depth=True)]
    M.backward_raw(st, r, *g3)
    assert st.num_rendered > 4 * tb["means3D"].shape[0] + 4096          # the regrow path ran
    assert guarded.check() > n1
