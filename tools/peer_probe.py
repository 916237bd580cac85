"""Developer tool (torchrun, >= 2 GPUs): standalone time of the peer all-reduce kernel vs NCCL for the ZJU payload."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from moss_b200 import parallel
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
dev = torch.device("cuda", local); torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
for n in (1_220_000, 6_100_000):
    buf = parallel.PeerGradBuffer(n, dev)
    x = torch.randn(n, device=dev)
    def timeit(fn, iters=200):
        for _ in range(20): fn()
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters): fn()
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / iters * 1e3
    t_peer = timeit(lambda: buf.allreduce(n))
    t_nccl = timeit(lambda: dist.all_reduce(x))
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        buf.allreduce(n)
    torch.cuda.current_stream().wait_stream(s); torch.cuda.synchronize()
    with torch.cuda.graph(g):
        for _ in range(10): buf.allreduce(n)
    t_graph = timeit(g.replay, 50) / 10
    t_push = timeit(lambda: buf.push_allreduce(n))
    g2 = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g2):
        for _ in range(10): buf.push_allreduce(n)
    t_push_graph = timeit(g2.replay, 50) / 10
    buf.check()
    if rank == 0:
        print(f"world {world} numel {n} ({n*4/1e6:.1f} MB): pull {t_peer:.1f} us  pull(graph x10) {t_graph:.1f} us  push {t_push:.1f} us  "
              f"push(graph x10) {t_push_graph:.1f} us  nccl {t_nccl:.1f} us  pad {buf.handle.signal_pad_size} B  multicast {hex(buf.multicast_ptr)}", flush=True)
dist.destroy_process_group()
