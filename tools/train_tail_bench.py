"""Measurement of the two kernels right behind the backward: densification statistics
and the AdamW step over the six per-Gaussian parameter groups, for ZJU / MonoCap sized Gaussian sets.
ours = one launch each through the C ABI; reference = the reference's own torch code on the same GPU
(masked updates of train_ZJU.py:171-174 / scene/gaussian_model.py:816-818; torch.optim.AdamW(l, lr=0.0, eps=1e-15),
scene/gaussian_model.py:225, default implementation and torch's fused=True).  One JSON line per size.
    python tools/train_tail_bench.py [P ...]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import moss_b200 as M
from moss_b200 import _abi
from moss_b200.optim import FusedAdamW

dev = torch.device("cuda:0")
lib = _abi.load()
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
hbm = json.load(open(pk)).get("hbm_gbs", 6650.0) if os.path.exists(pk) else 6650.0
SHAPES = {"xyz": 3, "f_dc": 3, "f_rest": 45, "opacity": 1, "scaling": 3, "rotation": 4}
LRS = {"xyz": 8e-4, "f_dc": 2.5e-3, "f_rest": 1.25e-4, "opacity": 0.05, "scaling": 5e-3, "rotation": 1e-3}
flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=dev)


def gpu_time(step, n=50):
    for _ in range(5):
        step()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(n):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); step(); e1.record(); torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / n * 1e3       # us


def kernel_us(name, step, n=30):
    lib.moss_profile_enable(1)
    for _ in range(n):
        flush.zero_(); step()
    torch.cuda.synchronize()
    ms, cnt = _abi.profile_read()[name]
    lib.moss_profile_enable(0)
    return ms / cnt * 1e3


for P in [int(a) for a in sys.argv[1:]] or [20000, 100000]:
    g = torch.Generator().manual_seed(P)
    radii = (torch.rand(P, generator=g) * 20 - 4).clamp_min(0).to(torch.int32).to(dev)      # ~20 % culled
    vgrad = torch.randn(P, 3, generator=g).to(dev)
    acc, den, mx = torch.zeros(P, 1, device=dev), torch.zeros(P, 1, device=dev), torch.zeros(P, device=dev)

    def stats_ref():
        vis = radii > 0
        mx[vis] = torch.max(mx[vis], radii[vis])
        acc[vis] += torch.norm(vgrad[vis, :2], dim=-1, keepdim=True)
        den[vis] += 1
    stats_ours = lambda: M.add_densification_stats(radii, vgrad, acc, den, mx)   # noqa: E731
    t_sr, t_so = gpu_time(stats_ref, 20), gpu_time(stats_ours)
    k_s = kernel_us("densify_stats", stats_ours)

    def groups():
        return [{"params": [torch.nn.Parameter(torch.randn(P, c, generator=g).to(dev))], "lr": LRS[n], "name": n} for n, c in SHAPES.items()]
    opts = {"ours": FusedAdamW(groups(), lr=0.0, eps=1e-15), "torch_default": torch.optim.AdamW(groups(), lr=0.0, eps=1e-15),
            "torch_fused": torch.optim.AdamW(groups(), lr=0.0, eps=1e-15, fused=True)}
    t = {}
    for name, opt in opts.items():
        for grp in opt.param_groups:
            grp["params"][0].grad = torch.randn_like(grp["params"][0])
        t[name] = gpu_time(opt.step, 30)
    k_a = kernel_us("adamw_step", opts["ours"].step)
    n_el = P * sum(SHAPES.values())
    bytes_a, bytes_s = n_el * 28, P * (4 + 8 + 24)
    print(json.dumps({"metric": "post-backward tail (densification stats + AdamW step)", "gaussians": P, "parameters": n_el,
                      "densify_stats": {"ours_us": round(t_so, 2), "kernel_us": round(k_s, 2), "reference_torch_masked_ops_us": round(t_sr, 2),
                                        "roofline": {"bound": "hbm", "achieved": round(bytes_s / (k_s * 1e-6) / 1e9, 1), "peak": hbm, "unit": "GB/s",
                                                     "frac": round(bytes_s / (k_s * 1e-6) / 1e9 / hbm, 4)}},
                      "adamw_step": {"ours_us": round(t["ours"], 2), "kernel_us": round(k_a, 2), "torch_default_us": round(t["torch_default"], 2),
                                     "torch_fused_us": round(t["torch_fused"], 2), "launches": 1,
                                     "roofline": {"bound": "hbm", "achieved": round(bytes_a / (k_a * 1e-6) / 1e9, 1), "peak": hbm, "unit": "GB/s",
                                                  "frac": round(bytes_a / (k_a * 1e-6) / 1e9 / hbm, 4)}}}))
