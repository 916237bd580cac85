"""GPU, >= 2 devices: the peer-memory all-reduce kernel (moss_peer_allreduce) against NCCL on the same buffers, and
the data-parallel step (two ranks render different views; summed packed gradient == single-process sum)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dev = torch.device("cuda", rank)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        from moss_b200 import parallel
        n = 1_220_003                       # ZJU-sized packed gradient, deliberately not a multiple of 4
        buf = parallel.PeerGradBuffer(n, dev)
        for it in range(5):
            g = torch.Generator(device=dev).manual_seed(1000 * it + rank)
            x = torch.randn(n, device=dev, generator=g)
            buf.tensor[:n].copy_(x)
            ref = x.clone()
            dist.all_reduce(ref)
            buf.allreduce(n)
            torch.cuda.synchronize(dev)
            buf.check()
            if world == 2:
                assert torch.equal(buf.tensor[:n], ref), f"iteration {it}: peer all-reduce differs from NCCL"
            else:
                assert torch.allclose(buf.tensor[:n], ref, rtol=1e-5, atol=1e-5)
        # bit-identical on every rank (fixed summation order)
        mine = buf.tensor[:n].clone()
        other = mine.clone()
        dist.broadcast(other, src=0)
        assert torch.equal(mine, other)

        # data-parallel rasterizer step: each rank renders its own camera, gradients land in the symmetric buffer
        import moss_b200 as M
        from moss_b200 import synth
        from moss_b200.rasterizer import grad_layout
        from tests import scenes as S
        d = S.make_scene("small")
        cams = synth.ring_cameras(23, 128, 128)
        gc, gd, ga = (t.to(dev) for t in S.upstream_grads(128, 128))
        to = lambda t: torch.empty(0) if t is None else t.to(dev)  # noqa: E731
        total = grad_layout(d["means3D"].shape[0], 16, True, True)[3]
        pbuf = parallel.PeerGradBuffer(total, dev)

        def render(cam, out):
            rs = M.GaussianRasterizationSettings(**synth.settings_kwargs(cam, d["bg"], 3, 1.0, device=dev))
            c, r, dp, a, st = M.forward_raw(to(d["means3D"]), to(d["shs"]), to(None), to(d["opacities"]), to(d["scales"]),
                                            to(d["rotations"]), to(None), rs)
            M.backward_raw(st, r, gc, gd, ga, flags=M._abi.FLAG_DETERMINISTIC, out=out)
            return st
        st = render(cams[rank], pbuf.tensor)
        pbuf.allreduce(st.param_grad_numel)
        torch.cuda.synchronize(dev)
        pbuf.check()
        expect = torch.zeros(total, device=dev)
        for r in range(world):
            tmp = torch.empty(total, device=dev)
            render(cams[r], tmp)
            expect += tmp
        k = st.param_grad_numel
        err = float((pbuf.tensor[:k] - expect[:k]).abs().max()) / float(expect[:k].abs().max())
        assert err < 1e-6, err
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs (run with gpurun --gpus 2)")
def test_peer_allreduce_matches_nccl_and_dp_step():
    mp.spawn(_worker, args=(2, _free_port()), nprocs=2, join=True)
