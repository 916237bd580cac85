"""Front-end measurement: skinning + covariance forward+backward for ZJU / MonoCap sized
Gaussian sets.  ours = the fused sm_100a kernel pair through the C ABI; reference = the reference's python ops
(oracle/lbs_oracle.py, a line-by-line restatement pinned to the real functions) as torch kernels on the same GPU with
autograd, i.e. what render() runs today; cpu = the same restatement on the host cores.  Prints one JSON line each.
    python tools/skin_bench.py [P ...]"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from moss_b200 import skinning as K, synth, _abi
from oracle import lbs_oracle as L
from tests.lbs_cases import oracle_smpl

dev = torch.device("cuda:0")
lib = _abi.load()
body = synth.body()
_pk = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
peaks = json.load(open(_pk)) if os.path.exists(_pk) else {"hbm_gbs": 6650.0}


def _frame_tables_line():
    """Per-frame SMPL tables: ours = ONE launch forward + ONE backward; reference = the
    reference's python ops (oracle restatement) as torch kernels on the same GPU with autograd."""
    V = body.verts.shape[0]
    smpl = K.Smpl(body.verts.to(dev), body.shapedirs.to(dev), body.posedirs.t().reshape(V, 3, 207).contiguous().to(dev),
                  body.J_regressor.to(dev), body.weights.to(dev), body.parents.to(dev))
    osm = L.Smpl(*[t.to(dev) for t in oracle_smpl(body)])
    fr = synth.frame_params(7)
    params = {k: v.to(dev) for k, v in fr.params.items()}
    big = {k: v.to(dev) for k, v in fr.big_params.items()}
    cr = L.batch_rodrigues(0.05 * torch.randn(23, 3)).reshape(1, 23, 3, 3).to(dev).requires_grad_(True)
    params["poses"].requires_grad_(True)
    gA, gpo = torch.randn(24, 4, 4, device=dev), torch.randn(V, 3, device=dev)
    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=dev)
    cache = K.FrameCache()

    def ours():
        params["poses"].grad = None; cr.grad = None
        ft = K.frame_tables(smpl, params, big, cr, cache=cache)
        torch.autograd.backward([ft.A_tgt, ft.poseoff_tgt], [gA, gpo])

    def ref():
        params["poses"].grad = None; cr.grad = None
        ft = L.frame_tables(osm, params, big, cr)       # the reference recomputes the big-pose half every render too
        torch.autograd.backward([ft.A_tgt, ft.poseoff_tgt], [gA, gpo])

    def gpu_time(step, n, cold):
        for _ in range(5): step()
        torch.cuda.synchronize()
        tot = 0.0
        for _ in range(n):
            if cold: flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); step(); e1.record(); torch.cuda.synchronize()
            tot += e0.elapsed_time(e1)
        return tot / n
    out = {"metric": "per-frame SMPL tables fwd+bwd", "vertices": V}
    for cold in (True, False):
        tag = "cold_l2" if cold else "warm_l2"
        t_ours = gpu_time(ours, 100, cold)
        lib.moss_profile_enable(1)
        for _ in range(50):
            if cold: flush.zero_()
            ours()
        torch.cuda.synchronize()
        prof = _abi.profile_read(); lib.moss_profile_enable(0)
        kf = prof["smpl_frame_forward"][0] / prof["smpl_frame_forward"][1] * 1e3
        kb = prof["smpl_frame_backward"][0] / prof["smpl_frame_backward"][1] * 1e3
        # algorithmic bytes: posedirs 207*3V*4 (both ways), forward also shapedirs 3V*10*4 + J_regressor 24*V*4 + v_template,
        # writes 2 x 3V*4; backward reads d_poseoff 3V*4
        bytes_f = 4 * (207 * 3 * V + 30 * V + 24 * V + 3 * V + 6 * V)
        bytes_b = 4 * (207 * 3 * V + 3 * V)
        out[tag] = {"ours_ms": round(t_ours, 4), "kernel_us": {"smpl_frame_forward": round(kf, 2), "smpl_frame_backward": round(kb, 2)},
                    "launches_per_step": (prof["smpl_frame_forward"][1] + prof["smpl_frame_backward"][1]) / 50,
                    "roofline": {"bound": "hbm" if cold else "l2", "kernel": "smpl_frame_forward", "algorithmic_bytes": bytes_f,
                                 "achieved": round(bytes_f / (kf * 1e-6) / 1e9, 1), "peak": peaks["hbm_gbs"], "unit": "GB/s",
                                 "frac": round(bytes_f / (kf * 1e-6) / 1e9 / peaks["hbm_gbs"], 4)},
                    "roofline_backward": {"algorithmic_bytes": bytes_b, "achieved": round(bytes_b / (kb * 1e-6) / 1e9, 1),
                                          "frac": round(bytes_b / (kb * 1e-6) / 1e9 / peaks["hbm_gbs"], 4)}}
    t_ref = gpu_time(ref, 30, False)
    out["reference_torch_ops_on_gpu_ms"] = round(t_ref, 4)
    out["speedup_vs_reference_python_path"] = round(t_ref / out["warm_l2"]["ours_ms"], 1)
    print(json.dumps(out))


_frame_tables_line()
for P in [int(a) for a in sys.argv[1:]] or [20000, 100000]:
    g = synth.gaussians(P, body, seed=synth.SEED + 5)
    fr = synth.frame_params(7)
    ft_c = L.frame_tables(oracle_smpl(body), fr.params, fr.big_params)
    ft = K.FrameTables(*[t.to(dev).contiguous() for t in ft_c])
    gen = torch.Generator().manual_seed(P)
    delta = 0.3 * torch.randn(P, 24, generator=gen)
    gw, gc = torch.randn(P, 3, generator=gen), torch.randn(P, 6, generator=gen)
    leaf = lambda t, d: t.to(d).requires_grad_(True)  # noqa: E731
    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=dev)

    def make(d, fn):
        xyz, sc, rot, dl = leaf(g.xyz, d), leaf(g.scales, d), leaf(g.rotations, d), leaf(delta, d)
        vid, w = g.vert_ids.to(d), body.weights.to(d)
        gwd, gcd = gw.to(d), gc.to(d)
        def step():
            for t in (xyz, sc, rot, dl):
                t.grad = None
            m, cov = fn(xyz, vid, w, sc, rot, dl)
            torch.autograd.backward([m, cov], [gwd, gcd])
        return step
    ours = make(dev, lambda x, v, w, s, r, d: K.skin_and_covariance(x, v, w, ft, s, r, 1.0, d)[:2])
    def ref_fn(ftab):
        def f(x, v, w, s, r, d):
            sk, cov = L.skin_and_covariance(x, v.long(), w, ftab, s, r, 1.0, d)
            return sk.world_pts, cov
        return f
    ref = make(dev, ref_fn(L.FrameTables(*ft)))
    cpu = make("cpu", ref_fn(ft_c))

    def gpu_time(step, n=100):
        for _ in range(10): step()
        torch.cuda.synchronize()
        tot = 0.0
        for _ in range(n):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); step(); e1.record(); torch.cuda.synchronize()
            tot += e0.elapsed_time(e1)
        return tot / n
    l0 = lib.moss_kernel_launches()
    t_ours = gpu_time(ours)
    launches = (lib.moss_kernel_launches() - l0) / 110
    lib.moss_profile_enable(1)
    for _ in range(50):
        flush.zero_(); ours()
    torch.cuda.synchronize()
    prof = _abi.profile_read(); lib.moss_profile_enable(0)
    kf, kb = prof["skin_forward"][0] / prof["skin_forward"][1] * 1e3, prof["skin_backward"][0] / prof["skin_backward"][1] * 1e3
    t_ref = gpu_time(ref, 30)
    torch.set_num_threads(os.cpu_count() or 1)
    cpu(); t0 = time.perf_counter(); cpu(); t_cpu = (time.perf_counter() - t0) * 1e3
    # algorithmic bytes: fwd reads xyz 12 + vid 4 + weight row 96 + delta 96 + 3 offset rows 36 + scaling 12 + rot 16,
    # writes means3D 12 + cov 24 + transforms 36 + bweights 96; bwd reads the same inputs + 36 of upstream gradient,
    # writes d_xyz 12 + d_scaling 12 + d_rot 16 + d_delta 96
    bytes_f, bytes_b = P * (272 + 168), P * (272 + 36 + 136)
    print(json.dumps({"metric": "skin+covariance fwd+bwd MGaussians/s", "gaussians": P, "ours_ms": round(t_ours, 4),
                      "ours_launches_per_step": launches, "kernel_us": {"skin_forward": round(kf, 2), "skin_backward": round(kb, 2)},
                      "roofline": {"bound": "hbm", "kernel": "skin_backward", "achieved": round(bytes_b / (kb * 1e-6) / 1e9, 1),
                                   "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": round(bytes_b / (kb * 1e-6) / 1e9 / peaks["hbm_gbs"], 4)},
                      "value": round(P / 1e6 / (t_ours * 1e-3), 1), "reference_torch_ops_on_gpu_ms": round(t_ref, 4),
                      "speedup_vs_reference_python_path": round(t_ref / t_ours, 1),
                      "cpu_baseline": {"ms": round(t_cpu, 2), "cores": os.cpu_count(), "kind": "port"}}))

    # ---- nearest SMPL vertex: ours vs an exhaustive torch search on the same GPU vs the CPU oracle
    verts_d, xyz_d = body.verts.to(dev).contiguous(), g.xyz.to(dev).contiguous()
    V = verts_d.shape[0]
    near_ours = lambda: K.nearest_vertex(xyz_d, verts_d)                      # noqa: E731  (grid search, grid cached)
    near_exh = lambda: K.nearest_vertex(xyz_d, verts_d, grid=None)            # noqa: E731  (exhaustive kernel)
    near_torch = lambda: torch.cdist(xyz_d, verts_d).argmin(dim=1)            # noqa: E731  (fp32 on the GPU, P x V matrix in HBM)
    t_no, t_nt = gpu_time(near_ours, 50), gpu_time(near_torch, 20)
    lib.moss_profile_enable(1)
    for _ in range(20):
        near_exh()
    torch.cuda.synchronize()
    ke = _abi.profile_read()["nearest_vertex"]; lib.moss_profile_enable(0)
    assert torch.equal(near_ours(), near_exh())
    t0 = time.perf_counter(); ref_ids = L.nearest_vertex(body.verts, g.xyz); t_nc = (time.perf_counter() - t0) * 1e3
    agree = float((near_ours().cpu().long() == ref_ids).float().mean())
    lib.moss_profile_enable(1)
    for _ in range(20):
        near_ours()
    torch.cuda.synchronize()
    kn = _abi.profile_read()["nearest_vertex"]; lib.moss_profile_enable(0)
    kn_us = kn[0] / kn[1] * 1e3
    pairs = P * V
    print(json.dumps({"metric": "nearest-vertex Gpairs/s", "gaussians": P, "vertices": V, "ours_ms": round(t_no, 4),
                      "kernel_us": round(kn_us, 2), "exhaustive_kernel_us": round(ke[0] / ke[1] * 1e3, 2), "value": round(pairs / 1e9 / (kn_us * 1e-6), 1),
                      "note": "kernel_us / value: the grid search (value counts the P*V pairs the exhaustive search would examine); "
                              "roofline: the exhaustive kernel",
                      "roofline": {"bound": "issue", "note": "~9 thread instructions per candidate pair from shared memory; "
                                   "HBM traffic is P*16+V*12 bytes", "achieved_ginst_per_s": round(pairs * 9 / 32 / 1e9 / (ke[0] / ke[1] * 1e-3), 1),
                                   "peak_ginst_per_s": round(148 * 4 * 1.965, 1)},
                      "torch_cdist_argmin_on_gpu_ms": round(t_nt, 4), "agreement_with_fp64_oracle": agree,
                      "cpu_baseline": {"ms": round(t_nc, 2), "cores": os.cpu_count(), "kind": "port"}}))
