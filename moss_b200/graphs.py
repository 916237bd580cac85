"""CUDA-graph capture of one rasterizer training step (forward + backward [+ gradient exchange]).

At ZJU shape a step is eight short kernels; launched one by one from Python the GPU idles a few microseconds between
them and the host needs ~140 us per step, which becomes the bottleneck once several ranks share a host.  A captured
step replays with one launch.  Capture uses MOSS_FLAG_ASYNC (no host synchronisation inside the library): the binning
capacity is fixed at capture time from the instance count the same view produced before, and the overflow flag stays
on the device -- ``GraphedStep.check()`` reads it back after a run.
"""
import torch

from . import _abi
from . import rasterizer as R


class GraphedStep:
    """One view's forward+backward as a CUDA graph over STATIC input tensors.

    inputs: dict with the tensors of ``rasterizer.forward_raw`` (means3D, shs, colors_precomp, opacities, scales,
    rotations, cov3D_precomp; empty tensors for the absent ones).  ``raster_settings`` may be a LIST of settings: the
    step is then one batched call over those views (``rasterizer.forward_batch_raw``; upstream gradients [V,3,H,W] /
    [V,1,H,W]).  ``grad_out`` (optional) is the packed gradient buffer the backward writes; ``push`` (optional) a
    ``parallel.PeerGradBuffer``: the data-parallel exchange is fused into the captured backward; ``after`` (optional) is
    called inside the capture after the backward (e.g. a stand-alone gradient exchange)."""

    def __init__(self, inputs, raster_settings, grads, grad_out=None, after=None, flags=0, forward_only=False, push=None,
                 backward_flags=0):
        self.inputs, self.rs, self.grads = inputs, raster_settings, grads
        self.push, self.backward_flags = push, backward_flags      # parallel.PeerGradBuffer: exchange fused into the backward
        self.forward_only = forward_only      # eval render (render_ZJU.py:56-72 under no_grad): no backward in the graph
        self.flags = flags | _abi.FLAG_ASYNC
        dev = inputs["means3D"].device
        # learn the instance count of this view (synchronous call), then capture with a fixed capacity
        self._run(flags, grad_out, None)
        torch.cuda.synchronize(dev)
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(2):
                self._run(self.flags, grad_out, None)
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.out = self._run(self.flags, grad_out, after)
        self.state = self.out[4]

    def _run(self, flags, grad_out, after):
        i = self.inputs
        fwd = R.forward_batch_raw if isinstance(self.rs, list) else R.forward_raw
        color, radii, depth, alpha, st = fwd(i["means3D"], i["shs"], i["colors_precomp"], i["opacities"],
                                             i["scales"], i["rotations"], i["cov3D_precomp"], self.rs, flags=flags)
        g = None if self.forward_only else R.backward_raw(st, radii, *self.grads, flags=self.backward_flags, out=grad_out, push=self.push)
        if after is not None:
            after(st)
        return color, radii, depth, alpha, st, g

    def replay(self):
        self.graph.replay()
        return self.out

    def check(self):
        """Raise if the captured binning capacity was exceeded (outputs invalid)."""
        c = R.export_state_counters(self.state)
        if int(c[1]) != 0:
            raise RuntimeError(f"graphed step: binning capacity {int(c[2])} exceeded by num_rendered {int(c[0])}")
        return int(c[0])


class GraphedIteration:
    """A whole training iteration written against the PUBLIC API -- e.g. skinning.frame_tables -> skinning.nearest_vertex
    -> fused.skin_and_rasterize (or GaussianRasterizer) -> losses.image_losses -> ``loss.backward()`` ->
    add_densification_stats -- captured as ONE CUDA graph.  At ZJU size such an iteration is ~0.37 ms of kernels behind
    ~0.8 ms of Python; replayed it costs one launch.

    ``fn()`` runs the iteration over STATIC tensors (parameters, camera, pose, targets: copy new values into them between
    replays, as with any CUDA graph) and returns what the caller wants to read afterwards (loss, image ...).  It is called
    once synchronously (the library learns the instance count that sizes the binning workspace), twice on a side stream
    and once under capture, the last three inside ``rasterizer.async_launch()``.  Gradients: call
    ``optimizer.zero_grad(set_to_none=True)`` at the top of ``fn`` so that the ``.grad`` tensors are created inside the
    graph's memory pool; an optimizer whose step count lives on the host (FusedAdamW, torch's non-capturable ones) steps
    OUTSIDE the graph, after ``replay()``.  Keep no graph-attached tensor of an earlier eager iteration alive (store
    ``loss.detach()``, not ``loss``): such a graph pins the parameters' AccumulateGrad nodes to the stream they were
    created on -- the legacy default stream -- and the captured backward then fails with
    ``cudaErrorStreamCaptureImplicit``.
    ``check()`` raises if a replay needed more binning capacity than the capture reserved (1.25 x the instance count of
    the synchronous call plus 4096: capture again after densification or a large change of pose or camera)."""

    def __init__(self, fn, device=None):
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.device = dev
        with torch.cuda.device(dev):
            fn()
            torch.cuda.synchronize(dev)
            side = torch.cuda.Stream(device=dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side), R.async_launch():
                for _ in range(2):
                    fn()
            torch.cuda.current_stream(dev).wait_stream(side)
            torch.cuda.synchronize(dev)
            R.async_counters()                       # drop the warm-up runs' counters
            self.graph = torch.cuda.CUDAGraph()
            with R.async_launch(), torch.cuda.graph(self.graph):
                self.out = fn()
            self.counters = R.async_counters()       # written by every replay

    def replay(self):
        self.graph.replay()
        return self.out

    def check(self):
        """-> instance counts of the last replay's forwards; raises if one exceeded its captured capacity."""
        counts = []
        for c in self.counters:
            c = c.cpu()
            if int(c[1]) != 0:
                raise RuntimeError(f"graphed iteration: binning capacity {int(c[2])} exceeded by num_rendered {int(c[0])}; capture again")
            counts.append(int(c[0]))
        return counts
