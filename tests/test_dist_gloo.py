"""CPU, world_size 2, gloo: the N>1 host logic of the path -- view sharding, the packed parameter-gradient
all-reduce and the densification-stat reductions (moss_b200/parallel.py)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from moss_b200 import parallel
from moss_b200.rasterizer import grad_layout


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, P, M):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        lay, param_numel, zero_from, total = grad_layout(P, M, True, False)
        g = torch.Generator().manual_seed(100 + rank)
        buf = torch.randn(total, generator=g)
        local = buf.clone()
        parallel.allreduce_param_grads(buf, param_numel)
        # expected: sum of both ranks' prefixes; the blend-only tail (colors, conic) stays local
        exp = sum(torch.randn(total, generator=torch.Generator().manual_seed(100 + r)) for r in range(world))
        assert torch.allclose(buf[:param_numel], exp[:param_numel], atol=1e-6)
        assert torch.equal(buf[param_numel:], local[param_numel:])
        o, n = lay["means3D"]
        assert o == 0 and n == 3 * P and lay["opacity"][0] + lay["opacity"][1] <= param_numel

        # views: disjoint cover
        mine = parallel.shard_views(23, rank, world)
        allv = [None] * world
        dist.all_gather_object(allv, mine)
        flat = sorted(v for part in allv for v in part)
        assert flat == list(range(23))
        fm = parallel.shard_frames_major(23, 4, rank, world)
        allf = [None] * world
        dist.all_gather_object(allf, fm)
        assert sorted(x for part in allf for x in part) == [(f, c) for f in range(4) for c in range(23)]
        assert len({f for f, _ in fm}) == 2          # frames stay together on a rank

        # densification statistics
        radii = torch.tensor([0, 5, 3, 0], dtype=torch.int32) if rank == 0 else torch.tensor([2, 1, 0, 0], dtype=torch.int32)
        mx = torch.zeros(4); acc = torch.zeros(4, 1); den = torch.zeros(4, 1)
        gn = torch.full((4, 1), float(rank + 1))
        parallel.allreduce_densification_stats(radii, mx, gn, acc, den)
        assert mx.tolist() == [2.0, 5.0, 3.0, 0.0]
        assert acc.view(-1).tolist() == [2.0, 3.0, 1.0, 0.0]
        assert den.view(-1).tolist() == [1.0, 2.0, 1.0, 0.0]
    finally:
        dist.destroy_process_group()


def test_packed_gradient_allreduce_and_sharding_world2():
    mp.spawn(_worker, args=(2, _free_port(), 257, 16), nprocs=2, join=True)


def test_balanced_view_schedule():
    costs = [200.0, 185.0, 430.0, 176.0, 245.0, 190.0, 413.0, 181.0, 199.0]
    for world in (1, 2, 4, 8):
        sched = parallel.balanced_view_schedule(costs, world)
        assert all(len(s) == world for s in sched)
        assert all(len(set(s)) == min(world, len(costs)) for s in sched)          # distinct views inside a step
        for r in range(world):                                                     # every rank renders every view
            assert {s[r] for s in sched} == set(range(len(costs)))
        # the views of one step are neighbours in cost order
        order = sorted(range(len(costs)), key=lambda i: costs[i])
        pos = {v: k for k, v in enumerate(order)}
        for s in sched:
            ps = sorted(pos[v] for v in set(s))
            assert ps[-1] - ps[0] == len(ps) - 1
    import pytest
    with pytest.raises(ValueError):
        parallel.balanced_view_schedule([], 2)


def test_single_process_is_a_noop():
    buf = torch.arange(10.0)
    assert parallel.allreduce_param_grads(buf, 6) is buf and buf.tolist() == list(range(10))
    assert parallel.shard_views(5, 0, 1) == [0, 1, 2, 3, 4]
