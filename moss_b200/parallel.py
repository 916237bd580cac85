"""Data-parallel plumbing for the rasterizer path.

Views (camera x frame) are independent given replicated Gaussians and their parameter gradients add, so the only
exchange step of the path is ONE sum all-reduce over the packed parameter-gradient prefix that the backward writes
(``RasterState.grad_buf[:param_grad_numel]``, rasterizer.grad_layout).  One process per GPU (torchrun), NCCL over
NVLink on the GPU box; the same code runs over gloo on CPU tensors in tests/test_dist_gloo.py.
"""
import torch
import torch.distributed as dist


def shard_views(n_views, rank, world):
    """Indices of the views rank `rank` renders: round-robin over cameras, so every rank sees every part of the ring
    (the reference iterates a shuffled camera stack, train_ZJU.py:93-99)."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
    return list(range(rank, n_views, world))


def shard_frames_major(n_cams, n_frames, rank, world):
    """(frame, camera) pairs with frames kept together on a rank, so a rank reuses its per-frame skinning tables
    across its cameras.  Frames are dealt round-robin; leftover ranks split a frame's cameras."""
    pairs = [(f, c) for f in range(n_frames) for c in range(n_cams)]
    if n_frames >= world:
        return [(f, c) for f, c in pairs if f % world == rank]
    return pairs[rank::world]


def allreduce_param_grads(grad_buf, param_numel, group=None, average=False):
    """Sum (or average) the leading `param_numel` floats of the packed gradient buffer across ranks, in place."""
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return grad_buf
    view = grad_buf[:param_numel]
    dist.all_reduce(view, op=dist.ReduceOp.SUM, group=group)
    if average:
        view.div_(dist.get_world_size(group))
    return grad_buf


def allreduce_densification_stats(radii, max_radii2D, viewspace_grad_norm, xyz_gradient_accum, denom, group=None):
    """The training-side bookkeeping that consumes the rasterizer's outputs (scene/gaussian_model.py:816-818,
    train_ZJU.py:171-174) under data parallelism: max of radii, sum of the gradient-norm accumulator and of the
    visibility count.  All in place; returns nothing."""
    vis = radii > 0
    local_max = torch.where(vis, torch.maximum(max_radii2D, radii.to(max_radii2D.dtype)), max_radii2D)
    add = torch.where(vis[:, None], viewspace_grad_norm, torch.zeros_like(viewspace_grad_norm))
    cnt = vis[:, None].to(denom.dtype)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(local_max, op=dist.ReduceOp.MAX, group=group)
        dist.all_reduce(add, op=dist.ReduceOp.SUM, group=group)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM, group=group)
    max_radii2D.copy_(local_max)
    xyz_gradient_accum.add_(add)
    denom.add_(cnt)


class PeerGradBuffer:
    """Symmetric (peer-mapped) packed-gradient buffer for the data-parallel exchange over NVLink peer memory.

    Fused form (the default of bench.py): ``rasterizer.backward_raw(..., push=buf)`` -- the backward's producer kernel
    stores every gradient row into the staging area of the rank that owns it and one short kernel behind it gathers the
    sums (include/moss_splat.h: moss_rasterize_backward_push).
    Stand-alone form: write ``.tensor`` (``out=``), then ``allreduce(param_grad_numel)`` (moss_peer_allreduce: a
    two-shot all-reduce kernel over the mapped peer pointers) -- for gradients that were produced by something else.

    The allocation and the exchange of the IPC handles are plumbing (torch.distributed._symmetric_memory); the data
    path is this repo's kernel.  Raises RuntimeError when symmetric memory cannot be set up on this machine -- callers
    that want a library collective instead use ``allreduce_param_grads`` (NCCL)."""

    def __init__(self, numel, device, group=None):
        import ctypes as C
        from . import _abi
        import torch.distributed._symmetric_memory as symm
        if not dist.is_initialized():
            raise RuntimeError("PeerGradBuffer needs an initialised process group")
        self._lib = _abi.load()
        self._C = C
        group = group if group is not None else dist.group.WORLD
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        if self.world > 8:
            raise RuntimeError("peer all-reduce supports up to 8 ranks of one node")
        numel = (int(numel) + 3) // 4 * 4
        # one symmetric allocation: [packed gradient | staging area of the fused exchange: world slices of <= numel/world]
        self.slice_max = int(self._lib.moss_push_slice_floats(numel, self.world))
        try:
            self.storage = symm.empty(numel + self.world * self.slice_max, dtype=torch.float32, device=device)
            self.handle = symm.rendezvous(self.storage, group)
        except Exception as e:       # no IPC / fabric support on this box
            raise RuntimeError(f"symmetric memory unavailable: {e}") from e
        self.tensor = self.storage[:numel]
        self.stage = self.storage[numel:]
        self._desc = {}
        need = self._lib.moss_peer_signal_bytes(self.world)
        if int(self.handle.signal_pad_size) < need:
            raise RuntimeError(f"signal pad too small ({self.handle.signal_pad_size} < {need})")
        import os
        mc = 0
        # NVLS (reduction inside the switch) moves 1/world of the buffer per GPU; measured on B200 it ties with the
        # peer-pointer path at 4 and 8 ranks and loses at 2 (tools/peer_probe.py), so it is used from 4 ranks up
        want = os.environ.get("MOSS_PEER_MULTIMEM", "auto")
        if want == "1" or (want == "auto" and self.world >= 4):
            try:
                mc = int(getattr(self.handle, "multicast_ptr", 0) or 0)
            except Exception:
                mc = 0
        self.multicast_ptr = mc                 # 0: no NVLS mapping -> loads/stores on the peer pointers
        self.path = "nvls" if mc else "peer"
        self.kind = ("exchange fused into the backward (rows pushed to their owner's staging area over NVLink, one gather kernel, "
                     + ("multimem.st through the switch" if mc else "peer stores") + ")")
        self.error = torch.zeros(4, dtype=torch.int32, device=device)      # [0] timeout flag, [1] device-side epoch, [2] block counter
        self.device = device
        # everyone's pad is zero and mapped before the first kernel spins on it
        self.storage.zero_()
        torch.cuda.synchronize(device)
        dist.barrier(group)

    def raw_pointers(self):
        """(gradient buffers, staging areas, signal pads) of all ranks as mapped in this process (lists of ints)."""
        numel = self.tensor.numel()
        h = self.handle
        return ([int(p) for p in h.buffer_ptrs], [int(p) + 4 * numel for p in h.buffer_ptrs], [int(p) for p in h.signal_pad_ptrs])

    def descriptor(self, param_numel):
        """ctypes MossPush for a backward that sums the leading `param_numel` floats of ``tensor`` over the ranks
        (include/moss_splat.h: moss_rasterize_backward_push)."""
        d = self._desc.get(param_numel)
        if d is None:
            C = self._C
            from . import _abi
            if param_numel % 4 or not 0 < param_numel <= self.tensor.numel():
                raise RuntimeError(f"param_numel {param_numel} must be a positive multiple of 4 within the buffer")
            h = self.handle
            w = self.world
            numel = self.tensor.numel()
            grads = (C.c_void_p * w)(*[int(p) for p in h.buffer_ptrs])
            stages = (C.c_void_p * w)(*[int(p) + 4 * numel for p in h.buffer_ptrs])
            pads = (C.c_void_p * w)(*[int(p) for p in h.signal_pad_ptrs])
            d = _abi.MossPush()
            d.rank, d.world = self.rank, w
            d.grad_base = self.tensor.data_ptr()
            d.numel = int(param_numel)
            d.grad_ptrs, d.stage_ptrs, d.signal_ptrs = grads, stages, pads
            d.signal_pad_bytes = int(h.signal_pad_size)
            d.multicast_grad = self.multicast_ptr or None
            d.error_word_dev = self.error.data_ptr()
            d._keep = (grads, stages, pads)
            self._desc[param_numel] = d
        return d

    def allreduce(self, numel, offset=0):
        """Sum tensor[offset:offset+numel] over all ranks, in place, on the current stream (no host sync; the epoch
        lives on the device, so the call can be captured in a CUDA graph)."""
        from . import _abi
        st = self._lib.moss_peer_allreduce(_abi.stream_ptr(self.device), self.rank, self.world,
                                           int(self.handle.buffer_ptrs_dev), int(self.handle.signal_pad_ptrs_dev),
                                           int(self.handle.signal_pad_size), int(offset), int(numel), 0,
                                           self.error.data_ptr(), self.multicast_ptr or None)
        _abi.check(st, "moss_peer_allreduce")
        return self.tensor

    def push_allreduce(self, numel):
        """Sum tensor[:numel] (numel a multiple of 4) over all ranks, in place, with the push exchange (one kernel: rows
        pushed to their owners' staging areas, gathered sums stored to every rank; include/moss_splat.h:
        moss_push_allreduce).  Same result on every rank, summed in rank order."""
        import ctypes as C
        from . import _abi
        st = self._lib.moss_push_allreduce(_abi.stream_ptr(self.device), C.byref(self.descriptor(int(numel))))
        _abi.check(st, "moss_push_allreduce")
        return self.tensor

    def check(self):
        """Host-side check (synchronises): did any spin time out?"""
        if int(self.error[0].item()) != 0:
            raise RuntimeError("moss_peer_allreduce: a peer did not arrive (timeout)")


def balanced_view_schedule(costs, world):
    """Cost-balanced data-parallel schedule.  A step ends when its SLOWEST rank has finished (the gradient exchange is
    a barrier), and the views of one body differ by up to 40 % in cost (edge-on views stack thousands of splats in a
    few tiles).  So the views rendered in the same step should cost about the same: sort the views by cost and cut
    the order into consecutive groups of `world` (the last group is the `world` most expensive views, so it may
    overlap its neighbour instead of wrapping around to the cheapest ones).  One cycle = one step per group; from
    cycle to cycle the ranks rotate through the positions of a group, so over `world` cycles every rank has rendered
    every view.  -> list of steps, each a list of `world` view indices (len = groups * world)."""
    n = len(costs)
    if n == 0 or world <= 0:
        raise ValueError("need at least one view and one rank")
    order = sorted(range(n), key=lambda i: (costs[i], i))
    if n <= world:
        groups = [[order[r % n] for r in range(world)]]
    else:
        starts = list(range(0, n - world, world)) + [n - world]
        groups = [order[s:s + world] for s in starts]
    steps = []
    for cycle in range(world):
        for g in groups:
            steps.append([g[(r + cycle) % world] for r in range(world)])
    return steps
