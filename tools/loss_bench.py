"""Image-loss measurement: the fused kernel (ONE launch: L1 + L2 + SSIM + S3IM values and
gradients) against the reference's python ops (oracle/loss_oracle.py, a restatement pinned to utils/loss_utils.py) as
torch kernels on the same GPU with autograd, and on the host cores.  Prints one JSON line per image size.
    python tools/loss_bench.py [WxH ...]"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from moss_b200 import losses as ML, _abi
from oracle import loss_oracle as LO
from tests.golden.make_loss_golden import synth_case

dev = torch.device("cuda:0")
lib = _abi.load()
_pk = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
peaks = json.load(open(_pk)) if os.path.exists(_pk) else {"hbm_gbs": 6650.0}
flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=dev)
for size in sys.argv[1:] or ["512x512", "1024x1024", "1920x1080"]:
    W, H = (int(v) for v in size.split("x"))
    box = (W // 4 + 1, H // 9, W - W // 4 - 2, H - H // 7 - 1)          # a person-sized crop
    image, alpha, gt, bkgd, bound = (torch.from_numpy(a) for a in synth_case(W, H, 11, box))
    d = {k: v.to(dev) for k, v in dict(image=image, alpha=alpha, gt=gt, bkgd=bkgd, bound=bound).items()}
    bounds = ML.mask_bounds(d["bound"])
    rect = bounds[:5].tolist()

    def ours():
        ig, ag = d["image"].detach().requires_grad_(True), d["alpha"].detach().requires_grad_(True)
        loss, _ = ML.image_losses(ig, ag, d["gt"], d["bkgd"], d["bound"], bounds)
        loss.backward()

    def ref_fn(dd):
        def f():
            ig, ag = dd["image"].detach().requires_grad_(True), dd["alpha"].detach().requires_grad_(True)
            total, _ = LO.image_losses(ig, ag, dd["gt"], dd["bkgd"], dd["bound"])     # incl. the boundingRect host round trip
            total.backward()
        return f
    ref = ref_fn(d)
    cpu = ref_fn(dict(image=image, alpha=alpha, gt=gt, bkgd=bkgd, bound=bound))

    def gpu_time(step, n):
        for _ in range(5): step()
        torch.cuda.synchronize()
        tot = 0.0
        for _ in range(n):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); step(); e1.record(); torch.cuda.synchronize()
            tot += e0.elapsed_time(e1)
        return tot / n
    t_ours = gpu_time(ours, 100)
    lib.moss_profile_enable(1)
    for _ in range(50):
        flush.zero_(); ours()
    torch.cuda.synchronize()
    prof = _abi.profile_read(); lib.moss_profile_enable(0)
    k_us = prof["image_losses"][0] / prof["image_losses"][1] * 1e3
    t_ref = gpu_time(ref, 20)
    torch.set_num_threads(os.cpu_count() or 1)
    cpu(); t0 = time.perf_counter(); cpu(); t_cpu = (time.perf_counter() - t0) * 1e3
    # algorithmic bytes: image + gt (3 floats each) + alpha, bkgd, bound read once, dL/dimage (3) + dL/dalpha written
    bytes_alg = W * H * 4 * (3 + 3 + 3 + 3 + 1)
    print(json.dumps({"metric": "image losses fwd+bwd (L1 + L2 + SSIM + S3IM)", "image": [W, H], "crop": rect[:4], "ours_ms": round(t_ours, 4),
                      "kernel_us": round(k_us, 2), "launches_per_step": prof["image_losses"][1] / 50 + 1,
                      "roofline": {"bound": "issue (separable window passes in shared memory)", "algorithmic_bytes": bytes_alg,
                                   "achieved_gbs": round(bytes_alg / (k_us * 1e-6) / 1e9, 1), "hbm_frac": round(bytes_alg / (k_us * 1e-6) / 1e9 / peaks["hbm_gbs"], 4)},
                      "reference_torch_ops_on_gpu_ms": round(t_ref, 4), "speedup_vs_reference_python_path": round(t_ref / t_ours, 1),
                      "cpu_baseline": {"ms": round(t_cpu, 2), "cores": os.cpu_count(), "kind": "port"}}))
