#!/usr/bin/env python
"""bench.py -- fwd+bwd throughput of the differentiable Gaussian-splat rasterizer on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference|cpu] [--workload zju|monocap|stress]

A "step" is ONE training iteration of the reference (train_ZJU.py:100,133): rasterize one camera view forward
and backward over the whole Gaussian set.  Workload at N=1 = BASELINE.json configs[1]: ZJU-Mocap shape, 20 000
Gaussians, 512x512, SH degree 3, MOSS's default covariance path (cov3D_precomp).  With N>1 every rank renders its
own view of the same (replicated) Gaussians per step and the packed parameter gradient is all-reduced over NCCL
(weak scaling: views/step = N).

  value  : Mpix/s, inputs resident in HBM, CUDA-event timed per step; L2 is flushed between steps (outside the events).
  e2e    : same metric through the public API (GaussianRasterizer + autograd) with HOST buffers: every step copies
           the step's inputs from pinned host memory and reads image + gradients back.  `e2e.value` blocks the host on
           every call (forward outputs are read back on a side stream during the backward); `e2e.pipelined` keeps two
           calls in flight.  Both arms run the same two loops (e2e_step / e2e_pipelined).
  roofline / cpu_baseline / clocks / gpu_launches: see DESIGN.md "Measurement".

--impl reference : the UNMODIFIED reference CUDA rasterizer (oracle/_ref/diff_gaussian_rasterization, rebuilt for
                   sm_100) through its own Python API on the same inputs, same e2e protocol (the reference has no CPU path).
--impl cpu       : the CPU oracle port (pure PyTorch restatement) on the host cores, bounded sample.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

WORKLOADS = {
    "zju": dict(scene="zju", desc="ZJU-Mocap shape: 20000 Gaussians, 512x512, SH deg 3, cov3D_precomp path, fwd+bwd, 1 view/step/GPU"),
    "monocap": dict(scene="monocap", desc="MonoCap shape: 100000 Gaussians, 1024x1024, SH deg 3, cov3D_precomp path, fwd+bwd"),
    "stress": dict(scene="stress", desc="Stress: 1M Gaussians generic 3DGS scene, 1920x1080, SH deg 3, scale/rotation path, fwd+bwd"),
}


# ---------------------------------------------------------------------------------------------------------
def clock_sampler(dev_index, stop, out):
    """nvidia-smi sampler thread (subprocess; does not hold the GIL while waiting)."""
    q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    try:
        p = subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "50",
                              "-i", str(dev_index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
    except Exception:
        return
    def reader():
        for line in p.stdout:
            out.append((time.perf_counter(), line.strip()))
    t = threading.Thread(target=reader, daemon=True)
    t.start()
    stop.wait()
    p.terminate()
    t.join(timeout=1.0)


def summarize_clocks(samples, t0, t1):
    sm, mx, reasons = [], 0, set()
    names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
    for ts, line in samples:
        if ts < t0 or ts > t1:
            continue
        f = [x.strip() for x in line.split(",")]
        try:
            sm.append(float(f[0])); mx = max(mx, float(f[1]))
        except Exception:
            continue
        for n, v in zip(names, f[3:7]):
            if v.lower().startswith("active"):
                reasons.add(n)
    if not sm:
        return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
    sm.sort()
    return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


def measured_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0}, "fallback"


# ---------------------------------------------------------------------------------------------------------
def build_inputs(workload, rank, world):
    from tests import scenes as S
    from moss_b200 import synth
    d = S.make_scene(WORKLOADS[workload]["scene"])
    cam0 = d["cam"]
    if workload == "stress":
        cams = [cam0]
    else:
        cams = synth.ring_cameras(23, cam0.image_width, cam0.image_height)
    return d, cams


def algorithmic_bytes(P, V, R, N, T, M, D, has_sr):
    """SURVEY.md section 8(d) formula restated for THIS pipeline's kernels (bytes that must move at least once)."""
    cin = 28 if has_sr else 24
    b = {}
    b["preprocess_forward"] = P * (12 + 4 + cin + 12 * (D + 1) ** 2) + V * (4 + 8 + 16 + 16 + 8 + 1) + P * 8
    b["tile_scan"] = T * (4 + 8 + 4)
    b["scatter_keys"] = P * 4 + V * (8 + 4) + R * 8
    b["blend_forward"] = R * (8 + 4) + R * 40 + N * 28 + T * 8
    b["blend_backward"] = N * (20 + 8) + R * (4 + 40) + V * 44
    b["preprocess_backward"] = P * 4 + V * (12 + 24 + 44 + 12 * (D + 1) ** 2 + 16) + P * (12 + 24 + 12 * M + (28 if has_sr else 0))
    return b


def e2e_object(mpix, pipe_ms_step, sync_ms_step, h2d, d2h, steps, **extra):
    """`value`: one call at a time, the host blocked until the call's results are in host memory (per-step CUDA
    events).  `pipelined`: the same calls, copies and bytes with two calls in flight (e2e_pipelined)."""
    o = {"value": round(mpix / (sync_ms_step * 1e-3), 2), "unit": "Mpix/s", "ms_per_step": round(sync_ms_step, 5),
         "protocol": "synchronous call: pinned-host inputs -> device, fwd+bwd through the public API, image + gradients -> "
                     "pinned host (forward outputs read back on a side stream during the backward), host sync; L2 flushed "
                     "between steps outside the events",
         "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": steps,
         "pipelined": {"value": round(mpix / (pipe_ms_step * 1e-3), 2), "ms_per_step": round(pipe_ms_step, 5),
                       "note": "two calls in flight (upload / kernels / read-back of consecutive steps overlap), L2 flush inside "
                               "the loop; host-bound when the kernels are shorter than the Python call path"}}
    o.update(extra)
    return o


def e2e_pipelined(dev, host, lay_names, fwd_bwd, world, steps, warmup, flush, depth=2):
    """End-to-end throughput with `depth` calls in flight (both arms run exactly this loop).

    Every step still copies ITS inputs from pinned host memory and reads ITS image + gradients back to pinned host
    memory; what changes against the synchronous protocol is only that step i+1's upload (copy-in stream) and step
    i-1's read-back (copy-out stream) run beside step i's kernels.  `fwd_bwd(t, i, after_forward)` runs one
    forward+backward through the public API, calls `after_forward((color, depth, alpha, radii))` in between and
    returns the gradient tensors.  Device-timed over the whole loop (events on the compute stream, which joins the
    copy-out stream before the closing event); the L2 flush runs inside the loop.  Returns ms for `steps` steps."""
    main = torch.cuda.current_stream(dev)
    s_in, s_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    slots = [{"out": {}, "done": torch.cuda.Event(), "in": torch.cuda.Event(), "fwd": torch.cuda.Event(),
              "bwd": torch.cuda.Event(), "keep": None} for _ in range(depth)]

    def to_host(sl, j, r):
        if j not in sl["out"]:
            sl["out"][j] = torch.empty(r.shape, dtype=r.dtype, pin_memory=True)
        sl["out"][j].copy_(r.detach(), non_blocking=True)

    def one(i):
        sl = slots[i % depth]
        sl["done"].synchronize()                 # this slot's previous results are on the host; its buffers are free
        with torch.cuda.stream(s_in):
            t = {k: host[k].to(dev, non_blocking=True) for k in lay_names}
            sl["in"].record(s_in)
        main.wait_event(sl["in"])
        t = {k: v.requires_grad_(True) for k, v in t.items()}

        def after_forward(outs):
            sl["fwd"].record(main)
            s_out.wait_event(sl["fwd"])
            with torch.cuda.stream(s_out):
                for j, r in enumerate(outs):
                    to_host(sl, j, r)
            sl["keep_f"] = outs

        grads = fwd_bwd(t, i, after_forward)
        if world > 1:
            for g in grads:
                dist.all_reduce(g)
        sl["bwd"].record(main)
        s_out.wait_event(sl["bwd"])
        with torch.cuda.stream(s_out):
            for j, r in enumerate(grads):
                to_host(sl, 4 + j, r)
            sl["done"].record(s_out)
        sl["keep"] = (t, grads)                  # device buffers stay referenced until the slot's read-back is done

    def drain():
        for sl in slots:
            sl["done"].synchronize()
        torch.cuda.synchronize(dev)

    for i in range(warmup):
        one(i)
    drain()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(main)
    for i in range(steps):
        flush.zero_()
        one(i)
    main.wait_stream(s_out)
    e1.record(main)
    drain()
    if world > 1:
        dist.barrier()
    out_bytes = sum(v.numel() * v.element_size() for v in slots[0]["out"].values())
    return e0.elapsed_time(e1), out_bytes


def run_ours(args, rank, world, dev):
    import moss_b200 as M
    from moss_b200 import _abi, parallel, synth
    lib = _abi.load()
    d, cams = build_inputs(args.workload, rank, world)
    W, H = cams[0].image_width, cams[0].image_height
    P = d["means3D"].shape[0]
    N = W * H
    names = ("means3D", "shs", "colors_precomp", "opacities", "scales", "rotations", "cov3D_precomp")
    host = {k: (d[k].contiguous().pin_memory() if d[k] is not None else None) for k in names}
    devt = {k: (host[k].to(dev, non_blocking=True) if host[k] is not None else torch.empty(0)) for k in names}
    settings = [M.GaussianRasterizationSettings(**synth.settings_kwargs(c, d["bg"], d["sh_degree"], d["scale_modifier"], device=dev))
                for c in cams]
    from tests.scenes import upstream_grads
    gc, gd, ga = (t.to(dev) for t in upstream_grads(W, H, depth=False))
    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=dev)   # > 126 MB L2

    state = {}
    # N > 1: the packed parameter gradient is all-reduced by the library's own kernel over NVLink peer memory (the
    # backward writes straight into the symmetric buffer); NCCL only if symmetric memory cannot be set up here
    peer, collective_kind = None, "none"
    if world > 1:
        from moss_b200.rasterizer import grad_layout
        total = grad_layout(P, 16 if d["shs"] is not None else 0, d["shs"] is not None, d["scales"] is not None)[3]
        ok = torch.ones(1, device=dev)
        if os.environ.get("MOSS_COLLECTIVE", "peer") == "peer":
            try:
                peer = parallel.PeerGradBuffer(total, dev)
            except RuntimeError as e:
                print(f"[rank {rank}] peer all-reduce unavailable, using NCCL: {e}", file=sys.stderr)
                ok.zero_()
        else:
            ok.zero_()
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)          # all ranks or none
        if float(ok) == 0.0:
            peer = None
        collective_kind = "peer-memory kernel (moss_peer_allreduce)" if peer is not None else "nccl all_reduce"

    # Which view does this rank render at step i?  N = 1: the ring in order.  N > 1: a cost-balanced schedule
    # (parallel.balanced_view_schedule) from a short calibration pass, so that the ranks of one step wait for views of
    # similar cost; every rank still cycles through all views.
    sched = None

    def view_of(i):
        if os.environ.get("MOSS_SAME_VIEW"):        # developer probe: all ranks render the same view (no imbalance)
            return i % len(settings)
        if sched is None:
            return (i * world + rank) % len(settings)
        return sched[i % len(sched)][rank]

    # The timed device-resident loop replays one CUDA graph per view (fwd + bwd + gradient exchange captured once,
    # MOSS_FLAG_ASYNC); --no-graph, the per-kernel profiling pass and rank-local extra steps use stream launches.
    graphs = {}
    use_graph = (not args.no_graph) and (world == 1 or peer is not None)
    inputs = dict(devt)

    def graph_step(i):
        from moss_b200.graphs import GraphedStep
        v = view_of(i)
        gs = graphs.get(v)
        if gs is None:
            after = (lambda st_: peer.allreduce(st_.param_grad_numel)) if peer is not None else None
            gs = graphs[v] = GraphedStep(inputs, settings[v], (gc, gd, ga), grad_out=peer.tensor if peer is not None else None,
                                         after=after)
        gs.replay()
        state["st"], state["radii"] = gs.state, gs.out[1]
        return gs.state

    def step(i, collective=True):
        rs = settings[view_of(i)]
        color, radii, depth, alpha, st = M.forward_raw(devt["means3D"], devt["shs"], devt["colors_precomp"], devt["opacities"],
                                                       devt["scales"], devt["rotations"], devt["cov3D_precomp"], rs)
        M.backward_raw(st, radii, gc, gd, ga, out=peer.tensor if peer is not None else None)
        if world > 1 and collective:
            if peer is not None:
                peer.allreduce(st.param_grad_numel)
            else:
                parallel.allreduce_param_grads(st.grad_buf, st.param_grad_numel)
        state["st"], state["radii"] = st, radii
        return st

    def timed_loop(fn, steps, warmup):
        for i in range(warmup):
            fn(i)
        torch.cuda.synchronize(dev)
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        for i in range(steps):
            flush.zero_()                       # L2 flush, outside the timed events
            ev[i][0].record()
            fn(i)
            ev[i][1].record()
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        t1 = time.perf_counter()
        ms = sum(a.elapsed_time(b) for a, b in ev)
        return ms, t0, t1

    if world > 1 and len(settings) > 1:
        costs = torch.zeros(len(settings), device=dev, dtype=torch.float64)
        for v in range(len(settings)):
            rs_v = settings[v]
            evs = []
            for rep in range(4):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                _c, _r, _d, _a, _st = M.forward_raw(devt["means3D"], devt["shs"], devt["colors_precomp"], devt["opacities"],
                                                    devt["scales"], devt["rotations"], devt["cov3D_precomp"], rs_v)
                M.backward_raw(_st, _r, gc, gd, ga)
                e1.record()
                evs.append((e0, e1))
            torch.cuda.synchronize(dev)
            costs[v] = min(a.elapsed_time(b) for a, b in evs[1:])
        dist.all_reduce(costs, op=dist.ReduceOp.MAX)            # same numbers, hence the same schedule, on every rank
        sched = parallel.balanced_view_schedule(costs.tolist(), world)

    # clocks
    samples, stop = [], threading.Event()
    th = threading.Thread(target=clock_sampler, args=(dev.index, stop, samples), daemon=True)
    if rank == 0 and not os.environ.get("MOSS_NO_SAMPLER"):
        th.start()
        time.sleep(0.3)

    timed_step = step
    if use_graph:
        try:
            # capture every view this rank will replay BEFORE the timed loop; all ranks capture the same number of
            # graphs in the same order (each capture runs the exchange a few times)
            for i in range(len(sched) if sched is not None else len(settings)):
                graph_step(i)
            torch.cuda.synchronize(dev)
            timed_step = graph_step
        except Exception as e:                       # capture not possible here: say so and time the stream path
            if world > 1:
                raise
            print(f"[bench] CUDA-graph capture failed, timing stream launches: {e}", file=sys.stderr)
            use_graph = False
    ms_total, t0, t1 = timed_loop(timed_step, args.steps, args.warmup)
    if use_graph:
        for gs in graphs.values():
            gs.check()
    # keep the same load running until the sampler has seen it for >= 1 s (the timed region may be shorter than
    # nvidia-smi's sampling period); these extra steps are not part of any reported time
    t_ext = time.perf_counter()
    i = 0
    while rank == 0 and not args.quick and time.perf_counter() - t0 < 1.2 and time.perf_counter() - t_ext < 3.0:
        step(i, collective=False); i += 1      # rank 0 only: must not issue a collective the other ranks do not match
        if i % 64 == 0:
            torch.cuda.synchronize(dev)
    torch.cuda.synchronize(dev)
    t_load_end = time.perf_counter()

    # per-kernel times (same steps, events around every kernel launch inside the library, on the launching stream)
    lib.moss_profile_enable(1)
    launches0 = lib.moss_kernel_launches()
    for i in range(args.steps):
        flush.zero_()
        step(i)
    torch.cuda.synchronize(dev)
    launches_timed = lib.moss_kernel_launches() - launches0      # kernels of args.steps steps (a replayed graph runs the same ones)
    prof = _abi.profile_read()
    lib.moss_profile_enable(0)

    # ---- e2e: public API with host buffers -----------------------------------------------------------------
    rast = [M.GaussianRasterizer(rs) for rs in settings]
    lay_names = [k for k in names if host[k] is not None]
    out_host = {}

    side = torch.cuda.Stream(dev)          # read-back of the forward outputs runs beside the backward

    def to_host(j, r):
        if j not in out_host:
            out_host[j] = torch.empty(r.shape, dtype=r.dtype, pin_memory=True)
        out_host[j].copy_(r.detach(), non_blocking=True)   # detach: a pinned tensor must not join the autograd graph

    def fwd_bwd(t, i, after_forward):
        means2D = torch.zeros_like(t["means3D"], requires_grad=True)
        color, radii, depth, alpha = rast[view_of(i)](means3D=t["means3D"], means2D=means2D, shs=t.get("shs"),
                                                         colors_precomp=t.get("colors_precomp"), opacities=t["opacities"],
                                                         scales=t.get("scales"), rotations=t.get("rotations"),
                                                         cov3D_precomp=t.get("cov3D_precomp"))
        after_forward((color, depth, alpha, radii))
        torch.autograd.backward([color, alpha], [gc, ga])
        return [means2D.grad] + [t[k].grad for k in lay_names]

    def e2e_step(i):
        main = torch.cuda.current_stream(dev)
        t = {k: host[k].to(dev, non_blocking=True).requires_grad_(True) for k in lay_names}

        def after_forward(outs):
            side.wait_stream(main)
            with torch.cuda.stream(side):
                for j, r in enumerate(outs):
                    to_host(j, r)

        res = fwd_bwd(t, i, after_forward)
        if world > 1:
            for g in res:
                dist.all_reduce(g)
        for j, r in enumerate(res):
            to_host(4 + j, r)
        main.synchronize()                                 # the step's result is on the host
        side.synchronize()
        if os.environ.get("MOSS_TRACE_ALLOC"):
            ms_ = torch.cuda.memory_stats(dev)
            print("alloc-trace", i, ms_["num_device_alloc"], ms_["num_device_free"], ms_["reserved_bytes.all.current"] >> 20,
                  ms_["allocated_bytes.all.current"] >> 20, ms_["num_alloc_retries"], file=sys.stderr)

    e2e_steps = max(10, min(args.steps, 200)) if not args.quick else 2
    if os.environ.get("MOSS_HOST_PROFILE"):          # developer aid: where does the host spend the e2e step?
        import cProfile, pstats
        pr = cProfile.Profile(); pr.enable()
        timed_loop(e2e_step, e2e_steps, 3)
        pr.disable()
        pstats.Stats(pr, stream=sys.stderr).sort_stats("tottime").print_stats(32)
    def n_malloc():
        return torch.cuda.memory_stats(dev).get("num_device_alloc", 0)
    m0 = n_malloc()
    e2e_ms, _, _ = timed_loop(e2e_step, e2e_steps, max(3, min(args.warmup, 10)) if not args.quick else 1)
    e2e_mallocs = n_malloc() - m0
    pipe_ms, pipe_out_bytes = e2e_pipelined(dev, host, lay_names, fwd_bwd, world, e2e_steps, 3 if not args.quick else 1, flush)
    h2d = sum(host[k].numel() * 4 for k in lay_names)
    d2h = sum(v.numel() * v.element_size() for v in out_host.values())

    stop.set()
    if peer is not None:
        peer.check()
    st = state["st"]
    R = int(M.export_state(st, "counters")[0])
    V = int((state["radii"] > 0).sum())
    T = ((W + 15) // 16) * ((H + 15) // 16)

    # ---- reduce over ranks -----------------------------------------------------------------------------------
    ms_step = ms_total / args.steps
    e2e_ms_step = e2e_ms / e2e_steps
    pipe_ms_step = pipe_ms / e2e_steps
    if world > 1:
        tt = torch.tensor([ms_step, e2e_ms_step, pipe_ms_step], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms_step, e2e_ms_step, pipe_ms_step = float(tt[0]), float(tt[1]), float(tt[2])
    if rank != 0:
        return None
    mpix = world * N / 1e6
    peaks, peak_src = measured_peaks()
    ab = algorithmic_bytes(P, V, R, N, T, int(st.scene.M), int(st.scene.D), d["scales"] is not None)
    kern = {k: (v[0] / v[1] * 1e3 if v[1] else 0.0) for k, v in prof.items()}      # us per launch
    dom = max((k for k in ab), key=lambda k: kern.get(k, 0.0))
    dom_us = kern[dom]
    achieved = ab[dom] / (dom_us * 1e-6) / 1e9 if dom_us > 0 else 0.0
    total_us = sum(kern[k] for k in ab)
    roof = {"bound": "hbm", "kernel": dom, "achieved": round(achieved, 2), "peak": peaks["hbm_gbs"], "unit": "GB/s",
            "frac": round(achieved / peaks["hbm_gbs"], 5), "traffic": None, "peak_source": peak_src,
            "kernel_us": round(dom_us, 2), "algorithmic_bytes": ab[dom],
            "kernels_us": {k: round(v, 2) for k, v in kern.items() if v > 0 and k != "other"},
            "step_algorithmic_bytes": sum(ab.values()),
            "step_hbm_frac": round(sum(ab.values()) / (total_us * 1e-6) / 1e9 / peaks["hbm_gbs"], 5) if total_us else None,
            "note": "blend kernels are FP32/SFU-issue and latency bound at this shape (working set < L2); see DESIGN.md"}
    traffic_file = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if os.path.exists(traffic_file):
        try:
            roof["traffic"] = json.load(open(traffic_file)).get(args.workload, {}).get(dom)
        except Exception:
            pass
    facts_file = os.path.join(ROOT, "profiles", "ncu_kernel_facts.json")
    if os.path.exists(facts_file):
        try:      # what actually bounds the dominant kernel: issue slots and occupancy from the committed ncu capture
            roof["ncu"] = json.load(open(facts_file)).get(args.workload, {}).get(dom)
            if roof["ncu"] and dom_us > 0:
                # issue-slot roofline of the same kernel: warp instructions (ncu) / live duration vs 148 SMs x 4
                # schedulers x 1.965 GHz -- the ceiling that actually applies to the blend kernels
                peak_issue = 148 * 4 * 1.965e9
                ach = roof["ncu"]["warp_inst"] / (dom_us * 1e-6)
                roof["issue"] = {"achieved": round(ach / 1e9, 1), "peak": round(peak_issue / 1e9, 1), "unit": "Gwarp-inst/s",
                                 "frac": round(ach / peak_issue, 4)}
        except Exception:
            pass
    clocks = summarize_clocks(samples, t0, max(t1, t_load_end))
    line = {
        "metric": "fwd+bwd Mpix/s", "value": round(mpix / (ms_step * 1e-3), 2), "unit": "Mpix/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms_step, 5), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOADS[args.workload]["desc"], "gaussians": P, "visible": V, "instances": R,
                   "image": [W, H], "views_per_step": world, "l2": "flushed between steps (256 MiB fill, outside the timed events)",
                   "launch": "one CUDA graph per view (fwd+bwd+exchange captured once, MOSS_FLAG_ASYNC)" if use_graph else "stream launches",
                   "parallelism": f"dp{world} (views sharded, packed parameter gradient all-reduced: {collective_kind})" if world > 1 else "single GPU"},
        "e2e": e2e_object(mpix, pipe_ms_step, e2e_ms_step, h2d, d2h, e2e_steps, device_mallocs=e2e_mallocs),
        "gpu_launches": int(launches_timed),
        "roofline": roof, "clocks": clocks,
    }
    if world == 1 and not args.no_cpu_baseline and not args.quick:
        line["cpu_baseline"] = cpu_baseline(args.workload)
    return line


# ---------------------------------------------------------------------------------------------------------
def cpu_baseline(workload, budget_s=25.0):
    """The CPU oracle port (oracle/raster_oracle.py) on the host cores: one fwd+bwd view of the same workload
    (falls back to BASELINE.json configs[0], the 6 890-Gaussian ZJU view, when the full workload would exceed
    the budget)."""
    from oracle import raster_oracle as O
    from tests import scenes as S
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    name = WORKLOADS[workload]["scene"]
    sample = f"1 view fwd+bwd of the {workload} workload"
    if workload != "zju":
        name, sample = "zju", "1 view fwd+bwd of the zju workload (20000 Gaussians, 512x512): the full workload exceeds the CPU budget"
    d = S.make_scene(name)
    from moss_b200 import synth
    cam0 = d["cam"]
    cams = synth.ring_cameras(23, cam0.image_width, cam0.image_height)
    grads = S.upstream_grads(cam0.image_width, cam0.image_height, depth=False)
    t0 = time.perf_counter()
    views = 0
    for cam in cams:            # successive ring views until ~10 s of CPU work (bounded sample of the same workload)
        s = O.Settings(**synth.settings_kwargs(cam, d["bg"], d["sh_degree"], d["scale_modifier"]))
        fw = O.rasterize_forward(d["means3D"], d["shs"], d["colors_precomp"], d["opacities"], d["scales"], d["rotations"],
                                 d["cov3D_precomp"], s)
        O.rasterize_backward(fw, d["means3D"], d["shs"], d["colors_precomp"], d["scales"], d["rotations"],
                             d["cov3D_precomp"], s, *grads, dtype=torch.float32)
        views += 1
        if time.perf_counter() - t0 >= min(10.0, budget_s):
            break
    dt = time.perf_counter() - t0
    sample = sample.replace("1 view", f"{views} views")
    return {"value": round(views * cam0.image_width * cam0.image_height / 1e6 / dt, 4), "unit": "Mpix/s", "cores": cores,
            "kind": "port", "sample": sample, "seconds": round(dt, 2)}


def run_cpu_arm(args):
    cb = cpu_baseline(args.workload)
    return {"impl": "cpu", "metric": "fwd+bwd Mpix/s", "value": cb["value"], "unit": "Mpix/s", "n_gpus": args.gpus,
            "steps": 1, "warmup": 0, "ms_per_step": round(cb["seconds"] * 1e3, 2), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOADS[args.workload]["desc"]}, "cpu_baseline": cb,
            "e2e": {"value": cb["value"], "unit": "Mpix/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


# ---------------------------------------------------------------------------------------------------------
def run_reference(args, rank, world, dev):
    """The UNMODIFIED reference rasterizer through its own Python API (oracle/_ref install), same protocol."""
    ref_dir = os.path.join(ROOT, "oracle", "_ref")
    if not os.path.exists(os.path.join(ref_dir, "diff_gaussian_rasterization", "_C.so")):
        return {"impl": "reference", "unavailable": "oracle/_ref/diff_gaussian_rasterization/_C.so not built (python oracle/build_ref.py torch)"}
    sys.path.insert(0, ref_dir)
    import diff_gaussian_rasterization as DGR
    from moss_b200 import synth
    from tests.scenes import upstream_grads
    d, cams = build_inputs(args.workload, rank, world)
    W, H = cams[0].image_width, cams[0].image_height
    P = d["means3D"].shape[0]
    N = W * H
    names = ("means3D", "shs", "colors_precomp", "opacities", "scales", "rotations", "cov3D_precomp")
    host = {k: (d[k].contiguous().pin_memory() if d[k] is not None else None) for k in names}
    lay_names = [k for k in names if host[k] is not None]
    devt = {k: host[k].to(dev) for k in lay_names}
    settings = [DGR.GaussianRasterizationSettings(**synth.settings_kwargs(c, d["bg"], d["sh_degree"], d["scale_modifier"], device=dev))
                for c in cams]
    rast = [DGR.GaussianRasterizer(rs) for rs in settings]
    gc, gd, ga = (t.to(dev) for t in upstream_grads(W, H, depth=False))
    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=dev)
    out_host = {}

    side = torch.cuda.Stream(dev)          # same protocol as our arm: forward outputs are read back beside the backward

    def to_host(j, r):
        if j not in out_host:
            out_host[j] = torch.empty(r.shape, dtype=r.dtype, pin_memory=True)
        out_host[j].copy_(r.detach(), non_blocking=True)   # detach: a pinned tensor must not join the autograd graph

    def fwd_bwd(t, i, after_forward):
        means2D = torch.zeros_like(t["means3D"], requires_grad=True)
        color, radii, depth, alpha = rast[(i * world + rank) % len(rast)](means3D=t["means3D"], means2D=means2D, shs=t.get("shs"),
                                                         colors_precomp=t.get("colors_precomp"), opacities=t["opacities"],
                                                         scales=t.get("scales"), rotations=t.get("rotations"),
                                                         cov3D_precomp=t.get("cov3D_precomp"))
        after_forward((color, depth, alpha, radii))
        torch.autograd.backward([color, alpha], [gc, ga])
        return [means2D.grad] + [t[k].grad for k in lay_names]

    def dev_step(i):
        t = {k: devt[k].detach().requires_grad_(True) for k in lay_names}
        res = fwd_bwd(t, i, lambda outs: None)
        if world > 1:
            for g in res:
                dist.all_reduce(g)

    def e2e_step(i):
        main = torch.cuda.current_stream(dev)
        t = {k: host[k].to(dev, non_blocking=True).requires_grad_(True) for k in lay_names}

        def after_forward(outs):
            side.wait_stream(main)
            with torch.cuda.stream(side):
                for j, r in enumerate(outs):
                    to_host(j, r)

        res = fwd_bwd(t, i, after_forward)
        if world > 1:
            for g in res:
                dist.all_reduce(g)
        for j, r in enumerate(res):
            to_host(4 + j, r)
        main.synchronize()
        side.synchronize()

    def timed_loop(fn, steps, warmup):
        for i in range(warmup):
            fn(i)
        torch.cuda.synchronize(dev)
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)
        for i in range(steps):
            flush.zero_()
            ev[i][0].record()
            fn(i)
            ev[i][1].record()
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        return sum(a.elapsed_time(b) for a, b in ev)

    ms_step = timed_loop(dev_step, args.steps, args.warmup) / args.steps
    e2e_steps = max(10, min(args.steps, 200))
    e2e_ms_step = timed_loop(e2e_step, e2e_steps, max(3, min(args.warmup, 10))) / e2e_steps
    pipe_ms, _ = e2e_pipelined(dev, host, lay_names, fwd_bwd, world, e2e_steps, 3, flush)
    pipe_ms_step = pipe_ms / e2e_steps
    if world > 1:
        tt = torch.tensor([ms_step, e2e_ms_step, pipe_ms_step], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms_step, e2e_ms_step, pipe_ms_step = float(tt[0]), float(tt[1]), float(tt[2])
    if rank != 0:
        return None
    mpix = world * N / 1e6
    h2d = sum(host[k].numel() * 4 for k in lay_names)
    d2h = sum(v.numel() * v.element_size() for v in out_host.values())
    return {"impl": "reference", "metric": "fwd+bwd Mpix/s", "value": round(mpix / (ms_step * 1e-3), 2), "unit": "Mpix/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms_step, 5),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOADS[args.workload]["desc"], "gaussians": P, "image": [W, H], "views_per_step": world,
                       "reference": "unmodified DGR CUDA rasterizer rebuilt for sm_100 (the reference has no CPU path)",
                       "l2": "flushed between steps (256 MiB fill, outside the timed events)"},
            "e2e": e2e_object(mpix, pipe_ms_step, e2e_ms_step, h2d, d2h, e2e_steps)}


# ---------------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "cpu"])
    ap.add_argument("--workload", default="zju", choices=list(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="time stream launches instead of replaying captured CUDA graphs")
    ap.add_argument("--quick", action="store_true", help="profiling runs (ncu): no clock extension, 2 e2e steps, no CPU baseline")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "cpu":
        if rank == 0:
            print(json.dumps(run_cpu_arm(args)), flush=True)
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (the product has no CPU path); use --impl cpu for the CPU oracle arm")
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")     # stdout carries exactly one JSON line
        dist.init_process_group("nccl", device_id=dev)
    try:
        line = run_ours(args, rank, world, dev) if args.impl == "ours" else run_reference(args, rank, world, dev)
        if rank == 0 and line is not None:
            print(json.dumps(line), flush=True)
    finally:
        if world > 1:
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
