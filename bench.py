#!/usr/bin/env python
"""bench.py -- fwd+bwd throughput of the differentiable Gaussian-splat rasterizer on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference|cpu] [--workload zju|monocap|stress]
                    [--views-per-step V] [--forward-only]

A "step" is ONE training iteration of the reference (train_ZJU.py:100,133): rasterize one camera view forward
and backward over the whole Gaussian set.  Workload at N=1 = BASELINE.json configs[1]: ZJU-Mocap shape, 20 000
Gaussians, 512x512, SH degree 3, MOSS's default covariance path (cov3D_precomp).  With N>1 every rank renders its
own view of the same (replicated) Gaussians per step and the packed parameter gradient is summed over the ranks by the
library's own exchange kernels over NVLink peer memory (weak scaling: views/step = N).
--views-per-step V (default 1) makes a step ONE batched call over V cameras per GPU (BASELINE.json configs[3]:
23 ZJU cameras): the gradients of a rank's views are summed locally and exchanged once.

  value            : Mpix/s, inputs resident in HBM, the step replayed as one CUDA graph of raw C-ABI calls; per-step
                     CUDA events; L2 flushed between steps (outside the events).  mean (ms_per_step), median, p10, p90.
  value_stream     : the same raw C-ABI calls as stream launches (no graph).
  value_public_api : the same step through the drop-in API (GaussianRasterizer + autograd), device-resident inputs --
                     the loop the reference arm's `value` is measured with (like for like).
  e2e              : the public API with HOST buffers: every step copies the step's inputs from pinned host memory and
                     reads image + gradients back.  `e2e.value` blocks the host on every call (forward outputs are read
                     back on a side stream during the backward); `e2e.pipelined` keeps two calls in flight.  Both arms
                     run the same two loops (e2e_step / e2e_pipelined).
  config / run     : `config` names the workload and is the SAME dict in both arms (workload, gaussians, image, views, L2
                     protocol, parallelism); what is specific to an arm's run (instances, launch mode, collective) is in `run`.
  roofline / cpu_baseline / clocks / gpu_launches / extra_workloads / front_end / exchange_check: see DESIGN.md "Measurement".

--impl reference : the UNMODIFIED reference CUDA rasterizer (oracle/_ref/diff_gaussian_rasterization, rebuilt for
                   sm_100) through its own Python API on the same inputs, same e2e protocol (the reference has no CPU
                   path); `value_raw_abi` times its kernels through the C shim with preallocated buffers (no autograd).
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
NAMES = ("means3D", "shs", "colors_precomp", "opacities", "scales", "rotations", "cov3D_precomp")


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


def dist_stats(ms):
    """mean / median / p10 / p90 of a list of per-step milliseconds."""
    s = sorted(ms)
    n = len(s)
    q = lambda f: s[min(n - 1, max(0, int(round(f * (n - 1)))))]  # noqa: E731
    return {"mean": sum(s) / n, "median": q(0.5), "p10": q(0.1), "p90": q(0.9)}


# ---------------------------------------------------------------------------------------------------------
def build_inputs(workload):
    from tests import scenes as S
    from moss_b200 import synth
    d = S.make_scene(WORKLOADS[workload]["scene"])
    cam0 = d["cam"]
    cams = [cam0] if workload == "stress" else synth.ring_cameras(23, cam0.image_width, cam0.image_height)
    return d, cams


def algorithmic_bytes(P, V, R, N, T, M, D, has_sr):
    """SURVEY.md section 8(d) formula restated for THIS pipeline's kernels (bytes that must move at least once).
    P, V, R, N, T are totals over the views of a batched step (per-view state rows)."""
    cin = 28 if has_sr else 24
    b = {}
    b["preprocess_forward"] = P * (12 + 4 + cin + 12 * (D + 1) ** 2) + V * (4 + 8 + 16 + 16 + 8 + 1 + 48) + P * 8
    b["tile_scan"] = T * (32 + 4 + 8 + 4)
    b["scatter_keys"] = P * 4 + V * (8 + 4) + R * 8
    b["tile_sort"] = R * (8 + 4)
    b["blend_forward"] = R * 4 + R * 40 + N * 28 + T * 8
    b["blend_backward"] = N * (20 + 8) + R * (4 + 40) + V * 48
    b["preprocess_backward"] = P * 4 + V * (12 + 24 + 96 + 12 * (D + 1) ** 2 + 16) + P * (12 + 24 + 12 * M + (28 if has_sr else 0))
    return b


def e2e_object(mpix, pipe_ms_step, sync, h2d, d2h, steps, **extra):
    """`value`: one call at a time, the host blocked until the call's results are in host memory (per-step CUDA
    events).  `pipelined`: the same calls, copies and bytes with two calls in flight (e2e_pipelined)."""
    o = {"value": round(mpix / (sync["mean"] * 1e-3), 2), "unit": "Mpix/s", "ms_per_step": round(sync["mean"], 5),
         "ms_median": round(sync["median"], 5), "ms_p10": round(sync["p10"], 5), "ms_p90": round(sync["p90"], 5),
         "protocol": "synchronous call: pinned-host inputs -> device, the step through the public API, image + gradients -> "
                     "pinned host (forward outputs read back on a side stream during the backward), host sync; L2 flushed "
                     "between steps outside the events",
         "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": steps,
         "pipelined": {"value": round(mpix / (pipe_ms_step * 1e-3), 2), "ms_per_step": round(pipe_ms_step, 5),
                       "note": "two calls in flight (upload / kernels / read-back of consecutive steps overlap), L2 flush inside "
                               "the loop; host-bound when the kernels are shorter than the Python call path"}}
    o.update(extra)
    return o


def e2e_pipelined(dev, host, lay_names, fwd_bwd, reduce_grads, steps, warmup, flush, depth=2):
    """End-to-end throughput with `depth` calls in flight (both arms run exactly this loop).

    Every step still copies ITS inputs from pinned host memory and reads ITS image + gradients back to pinned host
    memory; what changes against the synchronous protocol is only that step i+1's upload (copy-in stream) and step
    i-1's read-back (copy-out stream) run beside step i's kernels.  `fwd_bwd(t, i, after_forward)` runs one step
    through the public API, calls `after_forward((color, depth, alpha, radii))` in between and returns the gradient
    tensors; `reduce_grads(grads)` is the arm's data-parallel exchange.  Device-timed over the whole loop (events on
    the compute stream, which joins the copy-out stream before the closing event); the L2 flush runs inside the loop.
    Returns ms for `steps` steps."""
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

        grads = reduce_grads(fwd_bwd(t, i, after_forward))
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

    world = dist.get_world_size() if dist.is_initialized() else 1
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
    return e0.elapsed_time(e1)


def timed_loop(fn, steps, warmup, flush, dev, world):
    """W untimed warm-up steps, then K steps, each bracketed by its own CUDA events on the launching stream; the L2
    flush (256 MiB fill) sits between the steps, outside the events.  -> (list of per-step ms, wall t0, wall t1)."""
    for i in range(warmup):
        fn(i)
    torch.cuda.synchronize(dev)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    for i in range(steps):
        flush.zero_()
        ev[i][0].record()
        fn(i)
        ev[i][1].record()
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    t1 = time.perf_counter()
    return [a.elapsed_time(b) for a, b in ev], t0, t1


def max_over_ranks(vals, dev, world):
    if world == 1:
        return list(vals)
    tt = torch.tensor(list(vals), device=dev, dtype=torch.float64)
    dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    return [float(x) for x in tt]


# ---------------------------------------------------------------------------------------------------------
class OursHarness:
    """Device-resident inputs, settings, upstream gradients and the step functions of our arm for one workload."""

    def __init__(self, workload, dev, rank, world, V, forward_only, peer_ok=True):
        import moss_b200 as M
        from moss_b200 import parallel, synth
        from moss_b200.rasterizer import grad_layout
        from tests.scenes import upstream_grads
        self.M, self.dev, self.rank, self.world, self.V, self.forward_only = M, dev, rank, world, V, forward_only
        d, cams = build_inputs(workload)
        self.d, self.cams = d, cams
        self.W, self.H = cams[0].image_width, cams[0].image_height
        self.P = d["means3D"].shape[0]
        self.N = self.W * self.H
        self.host = {k: (d[k].contiguous().pin_memory() if d[k] is not None else None) for k in NAMES}
        self.devt = {k: (self.host[k].to(dev, non_blocking=True) if self.host[k] is not None else torch.empty(0)) for k in NAMES}
        self.lay_names = [k for k in NAMES if self.host[k] is not None]
        self.settings = [M.GaussianRasterizationSettings(**synth.settings_kwargs(c, d["bg"], d["sh_degree"], d["scale_modifier"], device=dev))
                         for c in cams]
        if V == 1:
            self.gc, self.gd, self.ga = (t.to(dev) for t in upstream_grads(self.W, self.H, depth=False))
        else:
            gs = [upstream_grads(self.W, self.H, seed=7 + v, depth=False) for v in range(V)]
            self.gc, self.gd, self.ga = (torch.stack([g[i] for g in gs]).to(dev) for i in range(3))
        # groups of V views a step renders: group g = views g*V .. g*V+V-1 (mod the ring)
        nv = len(self.settings)
        self.groups = [[(g * V + k) % nv for k in range(V)] for g in range((nv + V - 1) // V)] if V > 1 else [[v] for v in range(nv)]
        self.sched = None
        self.state = {}
        self.graphs = {}
        has_sh, has_sr = d["shs"] is not None, d["scales"] is not None
        self.total = grad_layout(self.P, 16 if has_sh else 0, has_sh, has_sr, V)[3]
        self.peer, self.collective_kind = None, "none"
        if world > 1 and not forward_only:
            ok = torch.ones(1, device=dev)
            if os.environ.get("MOSS_COLLECTIVE", "peer") == "peer" and peer_ok:
                try:
                    self.peer = parallel.PeerGradBuffer(self.total, dev)
                except RuntimeError as e:
                    print(f"[rank {rank}] peer all-reduce unavailable, using NCCL: {e}", file=sys.stderr)
                    ok.zero_()
            else:
                ok.zero_()
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)          # all ranks or none
            if float(ok) == 0.0:
                self.peer = None
            self.collective_kind = self.peer.kind if self.peer is not None else "nccl all_reduce"
            # Two peer-memory exchanges exist: "push" (fused into the backward: rows stored into their owner's staging area,
            # one gather kernel) and "pull" (one kernel behind the backward: flag barrier, NVLS multimem.ld_reduce / peer
            # loads, multimem.st, flag barrier).  Both are two NVLink hops and two flag phases; which one is shorter depends
            # on the world size and the switch, so MOSS_EXCHANGE=auto times both on this job's payload and keeps the faster.
            self.exchange_mode, self.exchange_us = "push", None
            if self.peer is not None:
                want = os.environ.get("MOSS_EXCHANGE", "auto")
                if want == "auto":
                    n = grad_layout(self.P, 16 if has_sh else 0, has_sh, has_sr, V)[1]
                    t = torch.zeros(2, device=dev, dtype=torch.float64)
                    for k, fn in enumerate((lambda: self.peer.push_allreduce(n), lambda: self.peer.allreduce(n))):
                        for _ in range(10):
                            fn()
                        torch.cuda.synchronize(dev); dist.barrier(); torch.cuda.synchronize(dev)
                        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        e0.record()
                        for _ in range(50):
                            fn()
                        e1.record(); torch.cuda.synchronize(dev)
                        t[k] = e0.elapsed_time(e1) / 50 * 1e3
                    dist.all_reduce(t, op=dist.ReduceOp.MAX)          # the same numbers, hence the same choice, on every rank
                    self.exchange_us = {"push": round(float(t[0]), 2), "pull": round(float(t[1]), 2)}
                    want = "pull" if float(t[1]) < float(t[0]) else "push"
                self.exchange_mode = want
                if want == "pull":
                    self.collective_kind = ("one exchange kernel behind the backward (flag barrier, "
                                            + ("multimem.ld_reduce + multimem.st through the switch" if self.peer.multicast_ptr else "peer loads + stores")
                                            + ", flag barrier)")

    # which group does this rank render at step i?  N = 1: the ring in order.  N > 1: a cost-balanced schedule
    # (parallel.balanced_view_schedule) from a short calibration pass, so that the ranks of one step wait for views of
    # similar cost; every rank still cycles through all groups.
    def group_of(self, i):
        if os.environ.get("MOSS_SAME_VIEW"):        # developer probe: all ranks render the same view (no imbalance)
            return i % len(self.groups)
        if self.sched is None:
            return (i * self.world + self.rank) % len(self.groups)
        return self.sched[i % len(self.sched)][self.rank]

    def rs_of(self, g):
        grp = self.groups[g]
        return self.settings[grp[0]] if self.V == 1 else [self.settings[v] for v in grp]

    def forward(self, rs, flags=0):
        t = self.devt
        f = self.M.forward_raw if self.V == 1 else self.M.forward_batch_raw
        return f(t["means3D"], t["shs"], t["colors_precomp"], t["opacities"], t["scales"], t["rotations"], t["cov3D_precomp"], rs, flags=flags)

    def step(self, i, collective=True):
        """One step as stream launches of the raw C ABI."""
        color, radii, depth, alpha, st = self.forward(self.rs_of(self.group_of(i)))
        if not self.forward_only:
            if self.peer is not None and collective and self.exchange_mode == "push":
                self.M.backward_raw(st, radii, self.gc, self.gd, self.ga, push=self.peer)      # exchange fused into the backward
            elif self.peer is not None and collective:
                self.M.backward_raw(st, radii, self.gc, self.gd, self.ga, out=self.peer.tensor)
                self.peer.allreduce(st.param_grad_numel)
            else:
                self.M.backward_raw(st, radii, self.gc, self.gd, self.ga)
                if self.world > 1 and collective:
                    self.exchange(st)
        self.state["st"], self.state["radii"] = st, radii
        return st

    def exchange(self, st):
        from moss_b200 import parallel
        parallel.allreduce_param_grads(st.grad_buf, st.param_grad_numel)        # NCCL: only when peer memory is unavailable

    def graph_step(self, i):
        from moss_b200.graphs import GraphedStep
        g = self.group_of(i)
        gs = self.graphs.get(g)
        if gs is None:
            pull = self.peer is not None and self.exchange_mode == "pull"
            gs = self.graphs[g] = GraphedStep(self.devt, self.rs_of(g), (self.gc, self.gd, self.ga),
                                              push=None if pull else self.peer,
                                              grad_out=self.peer.tensor if pull else None,
                                              after=(lambda st: self.peer.allreduce(st.param_grad_numel)) if pull else None,
                                              forward_only=self.forward_only)
        gs.replay()
        self.state["st"], self.state["radii"] = gs.state, gs.out[1]
        return gs.state

    def calibrate(self):
        """Per-group cost (ms), identical on every rank -> the balanced schedule."""
        from moss_b200 import parallel
        costs = torch.zeros(len(self.groups), device=self.dev, dtype=torch.float64)
        for g in range(len(self.groups)):
            evs = []
            for rep in range(4):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                _c, _r, _d, _a, _st = self.forward(self.rs_of(g))
                if not self.forward_only:
                    self.M.backward_raw(_st, _r, self.gc, self.gd, self.ga)
                e1.record()
                evs.append((e0, e1))
            torch.cuda.synchronize(self.dev)
            costs[g] = min(a.elapsed_time(b) for a, b in evs[1:])
        dist.all_reduce(costs, op=dist.ReduceOp.MAX)            # same numbers, hence the same schedule, on every rank
        self.sched = parallel.balanced_view_schedule(costs.tolist(), self.world)

    def capture_all(self):
        for i in range(len(self.sched) if self.sched is not None else len(self.groups)):
            self.graph_step(i)
        torch.cuda.synchronize(self.dev)

    def check_graphs(self):
        for gs in self.graphs.values():
            gs.check()


def run_ours(args, rank, world, dev):
    import moss_b200 as M
    from moss_b200 import _abi
    lib = _abi.load()
    V = args.views_per_step
    fo = args.forward_only
    h = OursHarness(args.workload, dev, rank, world, V, fo)
    W, H, P, N = h.W, h.H, h.P, h.N
    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=dev)   # > 126 MB L2
    if world > 1 and len(h.groups) > 1:
        h.calibrate()

    # clocks
    samples, stop = [], threading.Event()
    th = threading.Thread(target=clock_sampler, args=(dev.index, stop, samples), daemon=True)
    if rank == 0 and not os.environ.get("MOSS_NO_SAMPLER"):
        th.start()
        time.sleep(0.3)

    # The timed device-resident loop replays one CUDA graph per group of views (fwd + bwd + gradient exchange captured
    # once, MOSS_FLAG_ASYNC); --no-graph, the per-kernel profiling pass and rank-local extra steps use stream launches.
    use_graph = (not args.no_graph) and (world == 1 or h.peer is not None or fo)
    timed_step = h.step
    if use_graph:
        try:
            # capture every group this rank will replay BEFORE the timed loop; all ranks capture the same number of
            # graphs in the same order (each capture runs the exchange a few times)
            h.capture_all()
            timed_step = h.graph_step
        except Exception as e:                       # capture not possible here: say so and time the stream path
            if world > 1:
                raise
            print(f"[bench] CUDA-graph capture failed, timing stream launches: {e}", file=sys.stderr)
            use_graph = False
    ms_list, t0, t1 = timed_loop(timed_step, args.steps, args.warmup, flush, dev, world)
    if use_graph:
        h.check_graphs()
    # keep the same load running until the sampler has seen it for >= 1 s (the timed region may be shorter than
    # nvidia-smi's sampling period); these extra steps are not part of any reported time
    t_ext = time.perf_counter()
    i = 0
    while rank == 0 and not args.quick and time.perf_counter() - t0 < 1.2 and time.perf_counter() - t_ext < 3.0:
        h.step(i, collective=False); i += 1      # rank 0 only: must not issue a collective the other ranks do not match
        if i % 64 == 0:
            torch.cuda.synchronize(dev)
    torch.cuda.synchronize(dev)
    t_load_end = time.perf_counter()

    # the same raw C-ABI step as stream launches (what --no-graph times)
    stream_steps = max(3, min(args.steps, 100)) if not args.quick else 2
    ms_stream = ms_list if not use_graph else timed_loop(h.step, stream_steps, 3 if not args.quick else 1, flush, dev, world)[0]

    # per-kernel times (same steps, events around every kernel launch inside the library, on the launching stream)
    lib.moss_profile_enable(1)
    launches0 = lib.moss_kernel_launches()
    for i in range(args.steps):
        flush.zero_()
        h.step(i)
    torch.cuda.synchronize(dev)
    launches_timed = lib.moss_kernel_launches() - launches0      # kernels of args.steps steps (a replayed graph runs the same ones)
    prof = _abi.profile_read()
    lib.moss_profile_enable(0)

    # ---- the public API: GaussianRasterizer / rasterize_batch + autograd -------------------------------------------------
    rast = [M.GaussianRasterizer(rs) for rs in h.settings]
    lay_names = h.lay_names
    host = h.host
    gc, ga = h.gc, h.ga

    def fwd_bwd(t, i, after_forward):
        g = h.group_of(i)
        if V == 1:
            means2D = torch.zeros_like(t["means3D"], requires_grad=not fo)
            outs = rast[h.groups[g][0]](means3D=t["means3D"], means2D=means2D, shs=t.get("shs"),
                                        colors_precomp=t.get("colors_precomp"), opacities=t["opacities"],
                                        scales=t.get("scales"), rotations=t.get("rotations"),
                                        cov3D_precomp=t.get("cov3D_precomp"))
        else:
            means2D = torch.zeros((V,) + tuple(t["means3D"].shape), device=dev, requires_grad=not fo)
            outs = M.rasterize_batch(t["means3D"], means2D, t["opacities"], h.rs_of(g), shs=t.get("shs"),
                                     colors_precomp=t.get("colors_precomp"), scales=t.get("scales"),
                                     rotations=t.get("rotations"), cov3D_precomp=t.get("cov3D_precomp"))
        color, radii, depth, alpha = outs
        after_forward((color, depth, alpha, radii))
        if fo:
            return []
        torch.autograd.backward([color, alpha], [gc, ga])
        return [means2D.grad] + [t[k].grad for k in lay_names]

    # data-parallel exchange of the public-API path: ONE exchange of the packed gradient.  The autograd Function hands
    # out views of one packed buffer (rasterizer.grad_layout), so the parameter gradients are flattened for free.
    if h.peer is not None:
        M.set_gradient_sink(h.peer)      # autograd's backward writes into the symmetric buffer with the exchange fused in

    def reduce_grads(grads):
        if world == 1 or fo or not grads:
            return grads
        if h.peer is not None:           # already summed over the ranks by the backward itself; the tensors are views of the
            return [g.clone() for g in grads]    # symmetric buffer, which the next backward overwrites: keep a copy
        flat = torch.cat([g.reshape(-1) for g in grads])
        dist.all_reduce(flat)
        out, o = [], 0
        for g in grads:
            out.append(flat[o:o + g.numel()].view(g.shape)); o += g.numel()
        return out

    grad_ctx = torch.no_grad() if fo else torch.enable_grad()

    def api_step(i):          # device-resident inputs: the loop the reference arm's `value` uses
        with grad_ctx:
            t = {k: h.devt[k].detach().requires_grad_(not fo) for k in lay_names}
            reduce_grads(fwd_bwd(t, i, lambda outs: None))

    api_steps = max(5, min(args.steps, 200)) if not args.quick else 2
    ms_api, _, _ = timed_loop(api_step, api_steps, max(3, min(args.warmup, 10)) if not args.quick else 1, flush, dev, world)

    # ---- e2e: public API with host buffers -----------------------------------------------------------------
    out_host = {}
    side = torch.cuda.Stream(dev)          # read-back of the forward outputs runs beside the backward

    def to_host(j, r):
        if j not in out_host:
            out_host[j] = torch.empty(r.shape, dtype=r.dtype, pin_memory=True)
        out_host[j].copy_(r.detach(), non_blocking=True)   # detach: a pinned tensor must not join the autograd graph

    def e2e_step(i):
        main = torch.cuda.current_stream(dev)
        with grad_ctx:
            t = {k: host[k].to(dev, non_blocking=True).requires_grad_(not fo) for k in lay_names}

            def after_forward(outs):
                side.wait_stream(main)
                with torch.cuda.stream(side):
                    for j, r in enumerate(outs):
                        to_host(j, r)

            res = reduce_grads(fwd_bwd(t, i, after_forward))
        for j, r in enumerate(res):
            to_host(4 + j, r)
        main.synchronize()                                 # the step's result is on the host
        side.synchronize()

    e2e_steps = max(10, min(args.steps, 200)) if not args.quick else 2
    if os.environ.get("MOSS_HOST_PROFILE"):          # developer aid: where does the host spend the e2e step?
        import cProfile, pstats
        pr = cProfile.Profile(); pr.enable()
        timed_loop(e2e_step, e2e_steps, 3, flush, dev, world)
        pr.disable()
        pstats.Stats(pr, stream=sys.stderr).sort_stats("tottime").print_stats(32)

    def n_malloc():
        return torch.cuda.memory_stats(dev).get("num_device_alloc", 0)
    m0 = n_malloc()
    ms_e2e, _, _ = timed_loop(e2e_step, e2e_steps, max(3, min(args.warmup, 10)) if not args.quick else 1, flush, dev, world)
    e2e_mallocs = n_malloc() - m0
    with grad_ctx:
        pipe_ms = e2e_pipelined(dev, host, lay_names, fwd_bwd, reduce_grads, e2e_steps, 3 if not args.quick else 1, flush)
    h2d = sum(host[k].numel() * 4 for k in lay_names)
    d2h = sum(v.numel() * v.element_size() for v in out_host.values())

    # ---- N > 1: is the exchange kernel's sum the sum?  (first timed group, NCCL on a clone as the yardstick) ----------
    exchange_check = None
    if world > 1 and not fo:
        det = _abi.FLAG_DETERMINISTIC          # deterministic local gradients, so that two backward runs hold the same bits
        color_, radii_, depth_, alpha_, st = h.forward(h.rs_of(h.group_of(0)))
        M.backward_raw(st, radii_, h.gc, h.gd, h.ga, flags=det)
        k = st.param_grad_numel
        ref = st.grad_buf[:k].clone()
        dist.all_reduce(ref)                   # the yardstick: NCCL on the rank-local gradient
        if h.peer is not None and h.exchange_mode == "push":
            M.backward_raw(st, radii_, h.gc, h.gd, h.ga, flags=det, push=h.peer)
        elif h.peer is not None:
            M.backward_raw(st, radii_, h.gc, h.gd, h.ga, flags=det, out=h.peer.tensor)
            h.peer.allreduce(st.param_grad_numel)
        else:
            h.exchange(st)
        torch.cuda.synchronize(dev)
        got = st.grad_buf[:k]
        err = float((got - ref).abs().max())
        scale = float(ref.abs().max())
        same = got.clone()
        dist.broadcast(same, src=0)
        ident = torch.tensor([1.0 if torch.equal(same, got) else 0.0], device=dev)
        dist.all_reduce(ident, op=dist.ReduceOp.MIN)
        exchange_check = {"max_abs_err": err, "max_abs_ref": scale, "rel": err / max(scale, 1e-30),
                          "bit_identical_to_nccl": bool(torch.equal(got, ref)),
                          "bit_identical_across_ranks": bool(float(ident) == 1.0),
                          "path": (("fused push + gather, " if h.exchange_mode == "push" else "pull kernel, ") + h.peer.path) if h.peer is not None else "nccl",
                          "standalone_us": getattr(h, "exchange_us", None), "numel": int(k)}

    stop.set()
    if h.peer is not None:
        h.peer.check()
    st = h.state["st"]
    R = int(M.export_state(st, "counters")[0])
    Vis = int((h.state["radii"] > 0).sum())
    T = ((W + 15) // 16) * ((H + 15) // 16)

    # ---- reduce over ranks -----------------------------------------------------------------------------------
    sd, ss, sa, se = dist_stats(ms_list), dist_stats(ms_stream), dist_stats(ms_api), dist_stats(ms_e2e)
    pipe_ms_step = pipe_ms / e2e_steps
    red = max_over_ranks([sd["mean"], sd["median"], sd["p10"], sd["p90"], ss["mean"], sa["mean"], sa["median"],
                          se["mean"], se["median"], se["p10"], se["p90"], pipe_ms_step], dev, world)
    (sd["mean"], sd["median"], sd["p10"], sd["p90"], ss["mean"], sa["mean"], sa["median"],
     se["mean"], se["median"], se["p10"], se["p90"], pipe_ms_step) = red
    if rank != 0:
        return None
    ms_step = sd["mean"]
    mpix = world * V * N / 1e6
    peaks, peak_src = measured_peaks()
    ab = algorithmic_bytes(P * V, Vis, R, N * V, T * V, int(st.scene.M), int(st.scene.D), h.d["scales"] is not None)
    if fo:
        ab = {k: v for k, v in ab.items() if k not in ("blend_backward", "preprocess_backward")}
    kern = {k: (v[0] / v[1] * 1e3 if v[1] else 0.0) for k, v in prof.items()}      # us per launch
    dom = max((k for k in ab), key=lambda k: kern.get(k, 0.0))
    dom_us = kern[dom]
    achieved = ab[dom] / (dom_us * 1e-6) / 1e9 if dom_us > 0 else 0.0
    total_us = sum(kern[k] for k in ab)
    hbm = peaks["hbm_gbs"]
    roof = {"bound": "hbm", "kernel": dom, "achieved": round(achieved, 2), "peak": hbm, "unit": "GB/s",
            "frac": round(achieved / hbm, 5), "traffic": None, "peak_source": peak_src,
            "kernel_us": round(dom_us, 2), "algorithmic_bytes": ab[dom],
            "kernels_us": {k: round(v, 2) for k, v in kern.items() if v > 0 and k != "other"},
            "kernels_us_note": "event brackets around every launch of a stream-launched pass; the two timestamp operations "
                               "of a bracket cost 2-10 us that the graph-replayed step does not pay (holding the stream until "
                               "the whole step is enqueued changes nothing, so it is not launch gaps); the ncu durations of "
                               "the same launches are in profiles/",
            "kernels_hbm_frac": {k: round(ab[k] / (kern[k] * 1e-6) / 1e9 / hbm, 4) for k in ab if kern.get(k, 0) > 0},
            "step_algorithmic_bytes": sum(ab.values()),
            "step_hbm_frac": round(sum(ab.values()) / (ms_step * 1e-3) / 1e9 / hbm, 5),
            "note": "blend kernels are FP32/SFU-issue and latency bound at this shape (working set < L2); see DESIGN.md"}
    wl_key = args.workload if V == 1 else f"{args.workload}_v{V}"
    for fname, key in (("ncu_traffic.json", "traffic"), ("ncu_kernel_facts.json", "ncu")):
        path = os.path.join(ROOT, "profiles", fname)
        if os.path.exists(path):
            try:
                roof[key] = json.load(open(path)).get(wl_key, {}).get(dom)
            except Exception:
                pass
    facts_all = {}
    try:
        facts_all = json.load(open(os.path.join(ROOT, "profiles", "ncu_kernel_facts.json"))).get(wl_key, {})
    except Exception:
        pass
    if roof.get("ncu") and dom_us > 0:
        # issue-slot roofline of the same kernel: warp instructions (ncu) / live duration vs 148 SMs x 4
        # schedulers x 1.965 GHz -- the ceiling that actually applies to the blend kernels
        peak_issue = 148 * 4 * 1.965e9
        ach = roof["ncu"]["warp_inst"] / (dom_us * 1e-6)
        roof["issue"] = {"achieved": round(ach / 1e9, 1), "peak": round(peak_issue / 1e9, 1), "unit": "Gwarp-inst/s",
                         "frac": round(ach / peak_issue, 4)}
    if facts_all.get("tile_sort") and kern.get("tile_sort", 0) > 0:
        # the per-tile sort has its own line: bytes against HBM and warp instructions against the issue rate
        peak_issue = 148 * 4 * 1.965e9
        ts_us = kern["tile_sort"]
        roof["tile_sort"] = {"kernel_us": round(ts_us, 2), "algorithmic_bytes": ab["tile_sort"],
                             "hbm_frac": round(ab["tile_sort"] / (ts_us * 1e-6) / 1e9 / hbm, 4),
                             "issue_frac": round(facts_all["tile_sort"]["warp_inst"] / (ts_us * 1e-6) / peak_issue, 4)}
    clocks = summarize_clocks(samples, t0, max(t1, t_load_end))
    metric = "fwd Mpix/s" if fo else "fwd+bwd Mpix/s"
    desc = WORKLOADS[args.workload]["desc"]
    if fo:
        desc = desc.replace("fwd+bwd", "forward only (eval render)")
    if V > 1:
        desc = desc.replace("1 view/step/GPU", f"{V} views/step/GPU in one batched call")
    line = {
        "metric": metric, "value": round(mpix / (ms_step * 1e-3), 2), "unit": "Mpix/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms_step, 5),
        "ms_median": round(sd["median"], 5), "ms_p10": round(sd["p10"], 5), "ms_p90": round(sd["p90"], 5),
        "ms_per_view": round(ms_step / V, 5),
        "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        # `config` names the workload and is the same dict in both arms; what is specific to this arm's run sits in `run`
        "config": {"workload": desc, "gaussians": P, "image": [W, H], "views_per_step": world * V, "views_per_gpu_per_step": V,
                   "l2": "flushed between steps (256 MiB fill, outside the timed events)",
                   "parallelism": f"dp{world}" if world > 1 else "single GPU"},
        "run": {"visible": Vis, "instances": R,
                "launch": "one CUDA graph per step (fwd+bwd+exchange captured once, MOSS_FLAG_ASYNC)" if use_graph else "stream launches",
                "collective": f"views sharded, packed parameter gradient summed over ranks: {h.collective_kind}" if world > 1 else "none"},
        "value_stream": {"value": round(mpix / (ss["mean"] * 1e-3), 2), "ms_per_step": round(ss["mean"], 5),
                         "note": "the same raw C-ABI calls as stream launches (no CUDA graph)"},
        "value_public_api": {"value": round(mpix / (sa["mean"] * 1e-3), 2), "ms_per_step": round(sa["mean"], 5),
                             "ms_median": round(sa["median"], 5), "steps": api_steps,
                             "binding": M.public_api_kind(),
                             "note": "device-resident inputs through GaussianRasterizer + autograd (stream launches): the loop the "
                                     "reference arm's `value` is measured with"},
        "e2e": e2e_object(mpix, pipe_ms_step, se, h2d, d2h, e2e_steps, device_mallocs=e2e_mallocs),
        "gpu_launches": int(launches_timed),
        "roofline": roof, "clocks": clocks,
    }
    if exchange_check is not None:
        line["exchange_check"] = exchange_check
    if world == 1 and not args.no_cpu_baseline and not args.quick:
        line["cpu_baseline"] = cpu_baseline(args.workload)
    return line


def extra_workloads(dev, steps=20, warmup=3):
    """MonoCap and Stress (BASELINE.json configs[2], [4]) with short step counts, fwd+bwd and -- for Stress -- forward
    only, so that the driver-recorded JSON line carries them: graph-replayed raw C-ABI steps, per-step events, L2 flushed."""
    out = {}
    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=dev)
    for wl, modes in (("monocap", (False,)), ("stress", (False, True))):
        for fo in modes:
            try:
                h = OursHarness(wl, dev, 0, 1, 1, fo)
                h.capture_all()
                ms, _, _ = timed_loop(h.graph_step, steps, warmup, flush, dev, 1)
                h.check_graphs()
                s = dist_stats(ms)
                mp = h.N / 1e6
                st = h.state["st"]
                import moss_b200 as M
                out[wl + ("_forward_only" if fo else "")] = {
                    "metric": "fwd Mpix/s" if fo else "fwd+bwd Mpix/s", "value": round(mp / (s["mean"] * 1e-3), 2),
                    "ms_per_step": round(s["mean"], 5), "ms_median": round(s["median"], 5), "ms_p10": round(s["p10"], 5),
                    "ms_p90": round(s["p90"], 5), "steps": steps, "warmup": warmup, "gaussians": h.P, "image": [h.W, h.H],
                    "instances": int(M.export_state(st, "counters")[0]), "views": len(h.groups)}
                del h
                torch.cuda.empty_cache()
            except Exception as e:        # never let an extra take the headline line down
                out[wl + ("_forward_only" if fo else "")] = {"error": f"{type(e).__name__}: {e}"}
    return out


def front_end_numbers(dev, iters=30):
    """The rows of SURVEY.md section 8 around the rasterizer, at ZJU size, so that the driver-recorded line carries them:
    per-frame SMPL tables (f1), image losses (f3), and the fused skin + rasterize call against the two-call form (NS-a).
    Kernel times from the library's own event brackets (stream launches, cold L2); every entry point is the public one."""
    import moss_b200 as M
    from moss_b200 import _abi, fused as F, losses as ML, skinning as K, synth
    from tests.scenes import upstream_grads
    lib = _abi.load()
    out = {}
    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=dev)

    def measure(step):
        for _ in range(5):
            step()
        torch.cuda.synchronize(dev)
        lib.moss_profile_enable(1)
        for _ in range(iters):
            flush.zero_(); step()
        torch.cuda.synchronize(dev)
        prof = _abi.profile_read(); lib.moss_profile_enable(0)
        return {k: round(v[0] / v[1] * 1e3, 2) for k, v in prof.items() if v[1]}
    try:
        body = synth.body()
        V = body.verts.shape[0]
        smpl = K.Smpl(body.verts.to(dev), body.shapedirs.to(dev), body.posedirs.t().reshape(V, 3, 207).contiguous().to(dev),
                      body.J_regressor.to(dev), body.weights.to(dev), body.parents.to(dev))
        fr = synth.frame_params(9)
        fr.params["Th"] = torch.zeros(1, 3); fr.params["R"] = torch.eye(3)
        params = {k: v.to(dev) for k, v in fr.params.items()}
        big = {k: v.to(dev) for k, v in fr.big_params.items()}
        params["poses"].requires_grad_(True)
        cache = K.FrameCache()
        gA, gpo = torch.randn(24, 4, 4, device=dev), torch.randn(V, 3, device=dev)

        def frame():
            params["poses"].grad = None
            ft = K.frame_tables(smpl, params, big, cache=cache)
            torch.autograd.backward([ft.A_tgt, ft.poseoff_tgt], [gA, gpo])
        u = measure(frame)
        out["smpl_frame_tables"] = {"kernel_us": {k: u[k] for k in ("smpl_frame_forward", "smpl_frame_backward")}, "launches": 2,
                                    "algorithmic_bytes_forward": 4 * (207 * 3 * V + 30 * V + 24 * V + 3 * V + 6 * V)}
        # fused skin + rasterize vs skin kernel -> HBM -> rasterizer (20 k Gaussians on the SMPL body, 512 x 512)
        P, W, H = 20000, 512, 512
        g = synth.gaussians(P, body, seed=synth.SEED + 9)
        with torch.no_grad():
            ft = K.frame_tables(smpl, params, big, cache=cache)
        cam = synth.ring_cameras(23, W, H)[2]
        rs = M.GaussianRasterizationSettings(**synth.settings_kwargs(cam, torch.zeros(3), 3, 1.0, device=dev))
        gc = upstream_grads(W, H, depth=False)[0].to(dev)
        leaf = lambda t: t.to(dev).clone().requires_grad_(True)  # noqa: E731
        xyz, sc, rot, shs, opac = leaf(g.xyz), leaf(g.scales), leaf(g.rotations), leaf(g.shs), leaf(g.opacities)
        vid, wts = g.vert_ids.to(dev), body.weights.to(dev)
        rast = M.GaussianRasterizer(rs)

        def two_call():
            for t in (xyz, sc, rot, shs, opac):
                t.grad = None
            m, cov, _, _, _ = K.skin_and_covariance(xyz, vid, wts, ft, sc, rot, 1.0)
            c, _, _, _ = rast(means3D=m, means2D=torch.zeros_like(m), shs=shs, colors_precomp=None, opacities=opac, scales=None,
                              rotations=None, cov3D_precomp=cov)
            c.backward(gc)

        def one_call():
            for t in (xyz, sc, rot, shs, opac):
                t.grad = None
            c, _, _, _, _ = F.skin_and_rasterize(rs, xyz, vid, wts, ft, sc, rot, opac, shs=shs, return_transforms=False)
            c.backward(gc)
        u2, u1 = measure(two_call), measure(one_call)
        out["fused_skin_rasterize"] = {
            "two_call_per_gaussian_kernels_us": round(u2["skin_forward"] + u2["preprocess_forward"] + u2["preprocess_backward"] + u2["skin_backward"], 2),
            "fused_per_gaussian_kernels_us": round(u1["preprocess_forward"] + u1["preprocess_backward"], 2),
            "two_call_launches": len(u2), "fused_launches": len(u1), "two_call_kernels_sum_us": round(sum(u2.values()), 2),
            "fused_kernels_sum_us": round(sum(u1.values()), 2), "hbm_bytes_not_moved": P * 36 * 4}
        # image losses on a render-sized pair with a person-sized crop
        from tests.golden.make_loss_golden import synth_case
        box = (W // 4 + 1, H // 9, W - W // 4 - 2, H - H // 7 - 1)
        image, alpha, gt, bkgd, bound = (torch.from_numpy(a).to(dev) for a in synth_case(W, H, 11, box))
        bounds = ML.mask_bounds(bound)

        def loss():
            ig, ag = image.detach().requires_grad_(True), alpha.detach().requires_grad_(True)
            l, _ = ML.image_losses(ig, ag, gt, bkgd, bound, bounds)
            l.backward()
        u = measure(loss)
        out["image_losses"] = {"kernel_us": u["image_losses"], "image": [W, H], "crop": bounds[:4].tolist(),
                               "terms": "masked L1 + mask L2 + SSIM + S3IM, values and dL/dimage, dL/dalpha in one launch"}
    except Exception as e:            # never let an extra take the headline line down
        out["error"] = f"{type(e).__name__}: {e}"
    return out


# ---------------------------------------------------------------------------------------------------------
def cpu_baseline(workload, budget_s=25.0):
    """The CPU oracle port on the host cores (reported baseline, not the target).
    `value`: oracle/raster_oracle.py, fwd+bwd ring views of the same workload until ~10 s of CPU work (falls back to the
    zju view when the full workload would exceed the budget).
    `config1`: BASELINE.json configs[0] as stated -- 6 890 SMPL-topology Gaussians: skinning (oracle/lbs_oracle.py, the
    restatement of coarse_deform_c2source / smplx lbs pinned to the reference by tests/golden/lbs_*.npz) + covariance
    build + SH degree-3 colour evaluation (the polynomial of utils/sh_utils.py::eval_sh, pinned by
    tests/golden/eval_sh.npz) + the raster oracle for one 512x512 camera, forward and fwd+bwd, best of 3."""
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
    out = {"value": round(views * cam0.image_width * cam0.image_height / 1e6 / dt, 4), "unit": "Mpix/s", "cores": cores,
           "kind": "port", "sample": sample, "seconds": round(dt, 2)}
    try:
        out["config1"] = cpu_config1(cores)
    except Exception as e:
        out["config1"] = {"error": f"{type(e).__name__}: {e}"}
    return out


def cpu_config1(cores):
    """BASELINE.json configs[0]: skinning of 6 890 SMPL-topology Gaussians + covariance + SH deg 3 + raster oracle,
    one 512x512 ZJU camera, PyTorch on the host cores."""
    from oracle import raster_oracle as O, lbs_oracle as L
    from moss_b200 import synth
    from tests.lbs_cases import oracle_smpl
    body = synth.body()
    g = synth.gaussians(synth.N_SMPL_VERTS, body)
    fr = synth.frame_params(0)
    smpl = oracle_smpl(body)
    cam = synth.ring_cameras(23, 512, 512)[0]
    bg = torch.zeros(3)
    s = O.Settings(**synth.settings_kwargs(cam, bg, 3, 1.0))
    W, H = 512, 512
    from tests.scenes import upstream_grads
    grads = upstream_grads(W, H, depth=False)
    big = body.verts                       # canonical (big-pose) surface the nearest-vertex search runs against
    t_lbs = t_sh = t_fwd = t_all = 1e9
    for rep in range(3):
        a = time.perf_counter()
        ft = L.frame_tables(smpl, fr.params, fr.big_params)              # Rodrigues, kinematic chain, blend-shape offsets
        vert_ids = L.nearest_vertex(big, g.xyz)                         # the per-render KNN (scene/gaussian_model.py:827)
        sk, cov3D = L.skin_and_covariance(g.xyz, vert_ids, body.weights, ft, g.scales, g.rotations, 1.0)
        b = time.perf_counter()
        dirs = sk.world_pts - cam.camera_center[None]
        colors = torch.clamp_min(O.sh_to_rgb(3, g.shs, dirs), 0.0)      # eval_sh + 0.5, clamp (gaussian_renderer/__init__.py:100-105)
        c = time.perf_counter()
        fw = O.rasterize_forward(sk.world_pts, None, colors, g.opacities, None, None, cov3D, s)
        e = time.perf_counter()
        O.rasterize_backward(fw, sk.world_pts, None, colors, None, None, cov3D, s, *grads, dtype=torch.float32)
        f = time.perf_counter()
        t_lbs, t_sh, t_fwd, t_all = min(t_lbs, b - a), min(t_sh, c - b), min(t_fwd, e - a), min(t_all, f - a)
    mp = W * H / 1e6
    return {"gaussians": int(g.xyz.shape[0]), "image": [W, H], "cores": cores,
            "skinning_covariance_ms": round(t_lbs * 1e3, 2), "sh_ms": round(t_sh * 1e3, 2),
            "forward_ms": round(t_fwd * 1e3, 2), "fwd_bwd_ms": round(t_all * 1e3, 2),
            "forward_mpix_s": round(mp / t_fwd, 4), "fwd_bwd_mpix_s": round(mp / t_all, 4), "timing": "best of 3, time.perf_counter"}


def run_cpu_arm(args):
    cb = cpu_baseline(args.workload)
    return {"impl": "cpu", "metric": "fwd+bwd Mpix/s", "value": cb["value"], "unit": "Mpix/s", "n_gpus": args.gpus,
            "steps": 1, "warmup": 0, "ms_per_step": round(cb["seconds"] * 1e3, 2), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOADS[args.workload]["desc"]}, "cpu_baseline": cb,
            "e2e": {"value": cb["value"], "unit": "Mpix/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


# ---------------------------------------------------------------------------------------------------------
def run_reference(args, rank, world, dev):
    """The UNMODIFIED reference rasterizer through its own Python API (oracle/_ref install), same protocol.
    --views-per-step V: the reference has no batched call, so a step is V consecutive single-view calls with the
    gradients accumulated by autograd (what a MOSS user would write)."""
    ref_dir = os.path.join(ROOT, "oracle", "_ref")
    if not os.path.exists(os.path.join(ref_dir, "diff_gaussian_rasterization", "_C.so")):
        return {"impl": "reference", "unavailable": "oracle/_ref/diff_gaussian_rasterization/_C.so not built (python oracle/build_ref.py torch)"}
    sys.path.insert(0, ref_dir)
    import diff_gaussian_rasterization as DGR
    from moss_b200 import synth
    from tests.scenes import upstream_grads
    V, fo = args.views_per_step, args.forward_only
    d, cams = build_inputs(args.workload)
    W, H = cams[0].image_width, cams[0].image_height
    P = d["means3D"].shape[0]
    N = W * H
    host = {k: (d[k].contiguous().pin_memory() if d[k] is not None else None) for k in NAMES}
    lay_names = [k for k in NAMES if host[k] is not None]
    devt = {k: host[k].to(dev) for k in lay_names}
    settings = [DGR.GaussianRasterizationSettings(**synth.settings_kwargs(c, d["bg"], d["sh_degree"], d["scale_modifier"], device=dev))
                for c in cams]
    rast = [DGR.GaussianRasterizer(rs) for rs in settings]
    gcs = [tuple(t.to(dev) for t in upstream_grads(W, H, seed=7 + (v if V > 1 else 0), depth=False)) for v in range(V)]
    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=dev)
    out_host = {}
    nv = len(rast)
    groups = [[(g * V + k) % nv for k in range(V)] for g in range((nv + V - 1) // V)] if V > 1 else [[v] for v in range(nv)]

    side = torch.cuda.Stream(dev)          # same protocol as our arm: forward outputs are read back beside the backward

    def to_host(j, r):
        if j not in out_host:
            out_host[j] = torch.empty(r.shape, dtype=r.dtype, pin_memory=True)
        out_host[j].copy_(r.detach(), non_blocking=True)   # detach: a pinned tensor must not join the autograd graph

    def fwd_bwd(t, i, after_forward):
        grp = groups[(i * world + rank) % len(groups)]
        m2s, cols, rads, deps, alps = [], [], [], [], []
        for k, v in enumerate(grp):
            means2D = torch.zeros_like(t["means3D"], requires_grad=not fo)
            color, radii, depth, alpha = rast[v](means3D=t["means3D"], means2D=means2D, shs=t.get("shs"),
                                                  colors_precomp=t.get("colors_precomp"), opacities=t["opacities"],
                                                  scales=t.get("scales"), rotations=t.get("rotations"),
                                                  cov3D_precomp=t.get("cov3D_precomp"))
            m2s.append(means2D); cols.append(color); rads.append(radii); deps.append(depth); alps.append(alpha)
        if V == 1:
            after_forward((cols[0], deps[0], alps[0], rads[0]))
        else:
            after_forward((torch.stack(cols), torch.stack(deps), torch.stack(alps), torch.stack(rads)))
        if fo:
            return []
        torch.autograd.backward([x for k in range(V) for x in (cols[k], alps[k])],
                                [x for k in range(V) for x in (gcs[k][0], gcs[k][2])])
        m2 = m2s[0].grad if V == 1 else torch.stack([m.grad for m in m2s])
        return [m2] + [t[k].grad for k in lay_names]

    def reduce_grads(res):
        if world > 1:
            for g in res:             # the reference has no packed gradient: one NCCL call per tensor
                dist.all_reduce(g)
        return res

    grad_ctx = torch.no_grad() if fo else torch.enable_grad()

    def dev_step(i):
        with grad_ctx:
            t = {k: devt[k].detach().requires_grad_(not fo) for k in lay_names}
            reduce_grads(fwd_bwd(t, i, lambda outs: None))

    def e2e_step(i):
        main = torch.cuda.current_stream(dev)
        with grad_ctx:
            t = {k: host[k].to(dev, non_blocking=True).requires_grad_(not fo) for k in lay_names}

            def after_forward(outs):
                side.wait_stream(main)
                with torch.cuda.stream(side):
                    for j, r in enumerate(outs):
                        to_host(j, r)

            res = reduce_grads(fwd_bwd(t, i, after_forward))
        for j, r in enumerate(res):
            to_host(4 + j, r)
        main.synchronize()
        side.synchronize()

    # kernel-level number: the reference's CUDA code through the C shim (oracle/ref_shim.cu) with preallocated buffers:
    # no autograd, no torch allocation; the nine gradient tensors are zero-filled per call as its glue does
    # (DGR/rasterize_points.cu:158-166).  Like for like with our arm's `value_stream` (raw C ABI, stream launches).
    raw = None
    try:
        from oracle import ref_shim
        if ref_shim.available() and world == 1:
            lib = ref_shim.load()
            calls = []
            for v in range(nv):
                dv = dict(d); dv["cam"] = cams[v]
                rc = ref_shim.RefCall(dv, device=dev)
                rc.forward()
                calls.append(rc)
            M_ = calls[0].M
            z = lambda *s: torch.zeros(s, device=dev)  # noqa: E731
            gbuf = dict(means2D=z(P, 3), conic=z(P, 2, 2), opacity=z(P, 1), colors=z(P, 3), means3D=z(P, 3), cov3D=z(P, 6),
                        sh=z(P, M_, 3), scales=z(P, 3), rotations=z(P, 4))
            outc, outd, outa = z(3, H, W), z(1, H, W), z(1, H, W)
            rad = torch.zeros(P, dtype=torch.int32, device=dev)
            gd0 = torch.zeros(1, H, W, device=dev)
            p_ = ref_shim._p

            def raw_step(i):
                for k, v in enumerate(groups[i % len(groups)]):
                    c = calls[v]
                    outc.zero_(); outd.zero_(); outa.zero_(); rad.zero_()          # torch::full(0) x4, DGR/rasterize_points.cu:68-71
                    R = lib.ref_forward(P, c.D, c.M, p_(c.bg), W, H, p_(c.means3D), p_(c.shs), p_(c.colors), p_(c.opac),
                                        p_(c.scales), c.mod, p_(c.rots), p_(c.cov), p_(c.view), p_(c.proj), p_(c.campos),
                                        c.tanfovx, c.tanfovy, 0, p_(outc), p_(outd), p_(outa), p_(rad), 0)
                    if fo:
                        continue
                    for g_ in gbuf.values():
                        g_.zero_()
                    lib.ref_backward(P, c.D, c.M, R, p_(c.bg), W, H, p_(c.means3D), p_(c.shs), p_(c.colors), p_(outa),
                                     p_(c.scales), c.mod, p_(c.rots), p_(c.cov), p_(c.view), p_(c.proj), p_(c.campos),
                                     c.tanfovx, c.tanfovy, p_(rad), p_(gcs[k][0]), p_(gd0), p_(gcs[k][2]), p_(gbuf["means2D"]),
                                     p_(gbuf["conic"]), p_(gbuf["opacity"]), p_(gbuf["colors"]), p_(gbuf["means3D"]),
                                     p_(gbuf["cov3D"]), p_(gbuf["sh"]), p_(gbuf["scales"]), p_(gbuf["rotations"]), 0)
            raw_steps = max(5, min(args.steps, 100))
            ms_raw, _, _ = timed_loop(raw_step, raw_steps, 3, flush, dev, 1)
            raw = dist_stats(ms_raw)
            raw["steps"] = raw_steps
    except Exception as e:
        raw = {"error": f"{type(e).__name__}: {e}"}

    ms_list, _, _ = timed_loop(dev_step, args.steps, args.warmup, flush, dev, world)
    e2e_steps = max(10, min(args.steps, 200))
    ms_e2e, _, _ = timed_loop(e2e_step, e2e_steps, max(3, min(args.warmup, 10)), flush, dev, world)
    with grad_ctx:
        pipe_ms = e2e_pipelined(dev, host, lay_names, fwd_bwd, reduce_grads, e2e_steps, 3, flush)
    pipe_ms_step = pipe_ms / e2e_steps
    sd, se = dist_stats(ms_list), dist_stats(ms_e2e)
    red = max_over_ranks([sd["mean"], sd["median"], sd["p10"], sd["p90"], se["mean"], se["median"], se["p10"], se["p90"],
                          pipe_ms_step], dev, world)
    sd["mean"], sd["median"], sd["p10"], sd["p90"], se["mean"], se["median"], se["p10"], se["p90"], pipe_ms_step = red
    if rank != 0:
        return None
    mpix = world * V * N / 1e6
    h2d = sum(host[k].numel() * 4 for k in lay_names)
    d2h = sum(v.numel() * v.element_size() for v in out_host.values())
    desc = WORKLOADS[args.workload]["desc"]
    if fo:
        desc = desc.replace("fwd+bwd", "forward only (eval render)")
    if V > 1:
        desc = desc.replace("1 view/step/GPU", f"{V} views/step/GPU (V single-view calls: the reference has no batched call)")
    line = {"impl": "reference", "metric": "fwd Mpix/s" if fo else "fwd+bwd Mpix/s", "value": round(mpix / (sd["mean"] * 1e-3), 2),
            "unit": "Mpix/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(sd["mean"], 5),
            "ms_median": round(sd["median"], 5), "ms_p10": round(sd["p10"], 5), "ms_p90": round(sd["p90"], 5),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": desc, "gaussians": P, "image": [W, H], "views_per_step": world * V, "views_per_gpu_per_step": V,
                       "l2": "flushed between steps (256 MiB fill, outside the timed events)",
                       "parallelism": f"dp{world}" if world > 1 else "single GPU"},
            "run": {"reference": "unmodified DGR CUDA rasterizer rebuilt for sm_100 (the reference has no CPU path)",
                    "launch": "stream launches through the reference's Python API",
                    "collective": "one NCCL all_reduce per gradient tensor: the reference has no packed gradient" if world > 1 else "none"},
            "e2e": e2e_object(mpix, pipe_ms_step, se, h2d, d2h, e2e_steps)}
    if raw is not None:
        if "error" in raw:
            line["value_raw_abi"] = raw
        else:
            line["value_raw_abi"] = {"value": round(V * N / 1e6 / (raw["mean"] * 1e-3), 2), "ms_per_step": round(raw["mean"], 5),
                                     "ms_median": round(raw["median"], 5), "steps": raw["steps"],
                                     "note": "the reference's CUDA code through the C shim with preallocated buffers (outputs and the nine "
                                             "gradient tensors zero-filled per call as its glue does): no autograd, no allocation"}
    return line


# ---------------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "cpu"])
    ap.add_argument("--workload", default="zju", choices=list(WORKLOADS))
    ap.add_argument("--views-per-step", type=int, default=1, help="cameras per GPU per step, rendered by ONE batched call (1..32)")
    ap.add_argument("--forward-only", action="store_true", help="eval render (render_ZJU.py:56-72 under no_grad): no backward")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the short MonoCap / Stress runs embedded as extra_workloads")
    ap.add_argument("--no-graph", action="store_true", help="time stream launches instead of replaying captured CUDA graphs")
    ap.add_argument("--quick", action="store_true", help="profiling runs (ncu): no clock extension, 2 e2e steps, no CPU baseline")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if not 1 <= args.views_per_step <= 32:
        raise SystemExit("--views-per-step must be 1..32")

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
        if (args.impl == "ours" and rank == 0 and line is not None and world == 1 and args.workload == "zju"
                and args.views_per_step == 1 and not args.forward_only and not args.quick and not args.no_extra):
            line["extra_workloads"] = extra_workloads(dev)
            line["front_end"] = front_end_numbers(dev)
        if rank == 0 and line is not None:
            print(json.dumps(line), flush=True)
    finally:
        if world > 1:
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
