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
