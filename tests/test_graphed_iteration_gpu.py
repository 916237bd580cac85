"""graphs.GraphedIteration: a whole training iteration written against the public API (frame tables -> nearest vertex ->
fused skin + rasterize -> image losses -> backward -> densification statistics) captured as one CUDA graph under
rasterizer.async_launch().  Replays must give what the eager calls give on the same static tensors, follow in-place
changes of those tensors, and report a binning overflow instead of rendering garbage silently."""
import pytest
import torch

import moss_b200 as M
from moss_b200 import fused as F, losses as ML, skinning as K, synth, rasterizer as R
from moss_b200.graphs import GraphedIteration
from oracle import lbs_oracle as L
from tests.golden.make_loss_golden import synth_case

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
P, W, H = 6000, 160, 128


class _Scene:
    def __init__(self, seed):
        dev = torch.device(DEV)
        body = synth.body()
        V = body.verts.shape[0]
        g = synth.gaussians(P, body, seed=synth.SEED + seed)
        fr = synth.frame_params(seed)
        fr.params["Th"] = torch.zeros(1, 3); fr.params["R"] = torch.eye(3)
        self.params = {k: v.to(dev) for k, v in fr.params.items()}
        self.big = {k: v.to(dev) for k, v in fr.big_params.items()}
        cam = synth.ring_cameras(23, W, H)[2]
        self.rs = M.GaussianRasterizationSettings(**synth.settings_kwargs(cam, torch.zeros(3), 3, 1.0, device=dev))
        box = (W // 4 + 1, H // 9, W - W // 4 - 2, H - H // 7 - 1)
        _, _, self.gt, self.bkgd, self.bound = (torch.from_numpy(a).to(dev) for a in synth_case(W, H, 11, box))
        self.bounds = ML.mask_bounds(self.bound)
        self.verts, self.wts = body.verts.to(dev), body.weights.to(dev)
        self.smpl = K.Smpl(self.verts, body.shapedirs.to(dev), body.posedirs.t().reshape(V, 3, 207).contiguous().to(dev),
                           body.J_regressor.to(dev), self.wts, body.parents.to(dev))
        self.g = g
        gen = torch.Generator().manual_seed(seed)
        self.cr0 = L.batch_rodrigues(0.05 * torch.randn(23, 3, generator=gen)).reshape(1, 23, 3, 3).to(dev)

    def leaves(self):
        leaf = lambda t: t.to(DEV).clone().requires_grad_(True)  # noqa: E731
        g = self.g
        return dict(xyz=leaf(g.xyz), scaling=leaf(g.scales), rotation=leaf(g.rotations), shs=leaf(g.shs), opacity=leaf(g.opacities),
                    poses=leaf(self.params["poses"]), cr=leaf(self.cr0))

    def iteration(self, t):
        cache = K.FrameCache()
        acc, den, mx = (torch.zeros(P, 1, device=DEV), torch.zeros(P, 1, device=DEV), torch.zeros(P, device=DEV))
        stats = (acc, den, mx)

        def fn():
            for x in t.values():
                x.grad = None
            pr = dict(self.params); pr["poses"] = t["poses"]
            ft = K.frame_tables(self.smpl, pr, self.big, t["cr"], cache=cache)
            vid = K.nearest_vertex(t["xyz"].detach(), self.verts)
            m2 = torch.zeros_like(t["xyz"], requires_grad=True)
            c, r, d, a, _ = F.skin_and_rasterize(self.rs, t["xyz"], vid, self.wts, ft, t["scaling"], t["rotation"], t["opacity"],
                                                means2D=m2, shs=t["shs"], return_transforms=False)
            loss, terms = ML.image_losses(c, a, self.gt, self.bkgd, self.bound, self.bounds)
            loss.backward()
            M.add_densification_stats(r, m2.grad, acc, den, mx)
            return loss.detach(), c.detach(), r
        return fn, stats


def _close(a, b, what, tol=2e-4):
    scale = float(b.abs().max()) + 1e-12
    err = float((a - b).abs().max()) / scale
    assert err <= tol, (what, err)


def test_replays_equal_eager_iterations_and_follow_their_static_inputs():
    sc = _Scene(3)
    te, tg = sc.leaves(), sc.leaves()
    eager, stats_e = sc.iteration(te)
    graphed, stats_g = sc.iteration(tg)
    gi = GraphedIteration(graphed, DEV)
    static_grads = {k: v.grad for k, v in tg.items()}            # the graph's own .grad tensors: every replay rewrites them
    assert all(v is not None for v in static_grads.values())
    for s in stats_g:
        s.zero_()                                                 # the capture procedure ran the iteration four times
    for step in range(3):
        if step:                                                  # new values into the SAME tensors, both sides
            with torch.no_grad():
                for k in ("xyz", "poses", "opacity"):
                    d = (0.004 if k == "xyz" else 0.02) * step * torch.randn_like(te[k])
                    te[k].add_(d); tg[k].add_(d)
        loss_e, img_e, radii_e = eager()
        loss_g, img_g, radii_g = gi.replay()
        torch.cuda.synchronize()
        assert gi.check()[0] > 1000
        assert torch.equal(radii_e, radii_g)
        assert torch.equal(img_e, img_g)                          # the forward is bit-reproducible
        assert abs(float(loss_e) - float(loss_g)) <= 1e-6 * max(1.0, abs(float(loss_e)))
        for k in te:
            _close(static_grads[k], te[k].grad, (step, k))
    for a, b, name in zip(stats_g, stats_e, ("xyz_gradient_accum", "denom", "max_radii2D")):
        _close(a, b, name)


def test_a_replay_that_outgrows_the_captured_capacity_is_reported():
    sc = _Scene(5)
    t = sc.leaves()
    fn, _ = sc.iteration(t)
    gi = GraphedIteration(fn, DEV)
    gi.replay(); torch.cuda.synchronize()
    n0 = gi.check()[0]
    with torch.no_grad():
        t["scaling"].mul_(4.0)                                    # four times larger splats: far more (tile, Gaussian) pairs
    gi.replay(); torch.cuda.synchronize()
    with pytest.raises(RuntimeError, match="capacity"):
        gi.check()
    with torch.no_grad():
        t["scaling"].div_(4.0)
    gi.replay(); torch.cuda.synchronize()
    assert gi.check()[0] == n0                                    # and the graph is fine again afterwards


def test_async_launch_is_off_outside_its_block_and_counters_are_collected():
    sc = _Scene(7)
    t = sc.leaves()
    fn, _ = sc.iteration(t)
    fn()                                                          # synchronous: learns the capacity
    assert R._default_flags == 0 and R.async_counters() == []
    with R.async_launch():
        assert R._default_flags & M._abi.FLAG_ASYNC
        loss_a, img_a, _ = fn()
    assert R._default_flags == 0
    (c,) = R.async_counters()
    c = c.cpu()
    assert int(c[0]) > 0 and int(c[1]) == 0 and int(c[2]) >= int(c[0])
    loss_s, img_s, _ = fn()
    assert torch.equal(img_a, img_s)
