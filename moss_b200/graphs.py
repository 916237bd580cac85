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
