"""GPU, >= 2 devices: the data-parallel exchange kernels against NCCL and against the single-process sum.

  * moss_peer_allreduce (stand-alone two-shot all-reduce over mapped peer memory) on ZJU-sized buffers;
  * the exchange FUSED into the backward (moss_rasterize_backward_push: producer kernels push gradient rows to their
    owner's staging area, one gather kernel): every rank renders its own camera(s); the summed packed gradient must
    equal the rank-ordered sum of the per-rank gradients computed in one process -- bit for bit with
    MOSS_FLAG_DETERMINISTIC (the gather adds the staged copies in rank order), and identically on every rank.
Parametrised over world in {2, 4, 8} (as many as the box has) and over the data phase: peer loads/stores
(MOSS_PEER_MULTIMEM=0) and NVLS multimem through the switch (=1).  Tolerances: the peer path has a fixed summation
order (bit-identical to NCCL at world 2, 1e-5 relative otherwise); the in-switch fp32 reduction of the stand-alone NVLS
path has no documented order: 1e-5 relative to NCCL, rank-to-rank identity is reported, not required.  The fused
exchange only uses multimem for the STORE of already-summed slices, so its bits do not depend on the data phase."""
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


def _worker(rank, world, port, multimem):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    os.environ["MOSS_PEER_MULTIMEM"] = "1" if multimem else "0"
    dev = torch.device("cuda", rank)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        from moss_b200 import parallel
        n = 1_220_003                       # ZJU-sized packed gradient, deliberately not a multiple of 4
        buf = parallel.PeerGradBuffer(n, dev)
        if multimem and not buf.multicast_ptr:
            if rank == 0:
                print("NVLS multicast mapping unavailable on this box: multimem variant not exercised")
            return
        for it in range(5):
            g = torch.Generator(device=dev).manual_seed(1000 * it + rank)
            x = torch.randn(n, device=dev, generator=g)
            buf.tensor[:n].copy_(x)
            ref = x.clone()
            dist.all_reduce(ref)
            buf.allreduce(n)
            torch.cuda.synchronize(dev)
            buf.check()
            if world == 2 and not multimem:
                assert torch.equal(buf.tensor[:n], ref), f"iteration {it}: peer all-reduce differs from NCCL"
            else:
                assert torch.allclose(buf.tensor[:n], ref, rtol=1e-5, atol=1e-5)
        # the push exchange on its own (moss_push_allreduce): rank-ordered sums, identical on every rank
        n4 = n // 4 * 4
        for it in range(3):
            g = torch.Generator(device=dev).manual_seed(77 * it + rank)
            x = torch.randn(n4, device=dev, generator=g)
            buf.tensor[:n4].copy_(x)
            ref = x.clone()
            dist.all_reduce(ref)
            buf.push_allreduce(n4)
            torch.cuda.synchronize(dev)
            buf.check()
            assert torch.allclose(buf.tensor[:n4], ref, rtol=1e-5, atol=1e-5)
            if world == 2:
                assert torch.equal(buf.tensor[:n4], ref)
            chk = buf.tensor[:n4].clone()
            dist.broadcast(chk, src=0)
            assert torch.equal(chk, buf.tensor[:n4]), "push exchange: ranks disagree"
        buf.tensor[:n].copy_(torch.randn(n, device=dev, generator=torch.Generator(device=dev).manual_seed(5 + rank)))
        buf.allreduce(n)
        mine = buf.tensor[:n].clone()
        other = mine.clone()
        dist.broadcast(other, src=0)
        same = torch.equal(mine, other)
        if not multimem:
            assert same                      # fixed summation order on the peer-pointer path
        elif rank == 1:
            print(f"world {world}: stand-alone NVLS all-reduce identical on ranks 0 and 1: {same}")

        # ---- the exchange fused into the backward -------------------------------------------------------------
        import moss_b200 as M
        from moss_b200 import synth, _abi
        from moss_b200.graphs import GraphedStep
        from moss_b200.rasterizer import grad_layout
        from tests import scenes as S
        for scene, W in (("small", 128), ("ragged", None)):
            d = S.make_scene(scene)
            W_, H_ = d["cam"].image_width, d["cam"].image_height
            cams = synth.ring_cameras(23, W_, H_)
            to = lambda t: torch.empty(0) if t is None else t.to(dev)  # noqa: E731
            inputs = [to(d[k]) for k in ("means3D", "shs", "colors_precomp", "opacities", "scales", "rotations", "cov3D_precomp")]
            P = d["means3D"].shape[0]
            for V in (1, 3):
                gs = [S.upstream_grads(W_, H_, seed=7 + v) for v in range(V)]
                if V == 1:
                    gc, gd, ga = (t.to(dev) for t in gs[0])
                else:
                    gc, gd, ga = (torch.stack([g[i] for g in gs]).to(dev) for i in range(3))
                total = grad_layout(P, 16, True, True, V)[3]
                pbuf = parallel.PeerGradBuffer(total, dev)

                def settings(r):
                    rs = [M.GaussianRasterizationSettings(**synth.settings_kwargs(cams[(r * V + k) % 23], d["bg"], d["sh_degree"], 1.0, device=dev))
                          for k in range(V)]
                    return rs[0] if V == 1 else rs

                def render(r, out=None, push=None):
                    fwd = M.forward_raw if V == 1 else M.forward_batch_raw
                    c, rad, dp, a, st = fwd(*inputs, settings(r))
                    M.backward_raw(st, rad, gc, gd, ga, flags=_abi.FLAG_DETERMINISTIC, out=out, push=push)
                    return st
                for rep in range(3):          # repeated exchanges reuse the staging area and the flag slots
                    st = render(rank, push=pbuf)
                torch.cuda.synchronize(dev)
                pbuf.check()
                k = st.param_grad_numel
                expect = None
                for r in range(world):
                    tmp = torch.zeros(total, device=dev)
                    render(r, out=tmp)
                    expect = tmp.clone() if expect is None else expect + tmp       # rank order, like the gather kernel
                got = pbuf.tensor[:k]
                assert torch.equal(got, expect[:k]), f"{scene} V={V}: fused exchange differs from the rank-ordered sum " \
                    f"(max abs {float((got - expect[:k]).abs().max())})"
                ref = expect[:k].clone()
                dist.broadcast(ref, src=0)
                assert torch.equal(got, ref), "ranks disagree"
                # the blend-only tail (colours, conics, per-view rows) stays local and untouched by the exchange
                mine_tail = torch.zeros(total, device=dev)
                render(rank, out=mine_tail)
                assert torch.equal(pbuf.tensor[k:total], mine_tail[k:total])
                # the same step captured in a CUDA graph and replayed
                ins = dict(zip(("means3D", "shs", "colors_precomp", "opacities", "scales", "rotations", "cov3D_precomp"), inputs))

                gstep = GraphedStep(ins, settings(rank), (gc, gd, ga), push=pbuf, backward_flags=_abi.FLAG_DETERMINISTIC)
                for rep in range(3):
                    pbuf.tensor.zero_()
                    gstep.replay()
                torch.cuda.synchronize(dev)
                pbuf.check()
                assert torch.equal(pbuf.tensor[:k], expect[:k]), f"{scene} V={V}: graph replay of the fused exchange differs"
                dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("multimem", [0, 1])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_exchange_kernels_match_nccl_and_the_single_process_sum(world, multimem):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs (run with gpurun --gpus {world})")
    mp.spawn(_worker, args=(world, _free_port(), multimem), nprocs=world, join=True)
