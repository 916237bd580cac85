"""Developer tool: fold one `ncu --set full` report into profiles/ncu_kernel_facts.json and profiles/ncu_traffic.json
(the two files bench.py reads for `roofline.ncu`, `roofline.issue` and `roofline.traffic`).
usage: python tools/ncu_facts.py report.ncu-rep WORKLOAD      (values are per launch: mean over the captured launches)"""
import collections, csv, io, json, os, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NAMES = ("preprocess_forward", "tile_scan", "scatter_keys", "tile_sort", "blend_forward", "blend_backward", "preprocess_backward",
         "skin_forward", "skin_backward", "peer_allreduce", "nearest_vertex", "densify_stats", "adamw", "sum_views", "smpl_frame_forward",
         "smpl_frame_backward", "image_losses")
FACTS = {"issue_active_pct": "smsp__issue_active.avg.pct_of_peak_sustained_active",
         "warps_active_pct": "sm__warps_active.avg.pct_of_peak_sustained_active",
         "pipe_fma_pct": "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
         "pipe_xu_pct": "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
         "warp_inst": "smsp__inst_executed.sum", "duration_us": "gpu__time_duration.sum"}
SCALE = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def main():
    rep, workload = sys.argv[1], sys.argv[2]
    # a report, or its `ncu -i report --page raw --csv` export (what travels back from the GPU box when reports are large)
    out = open(rep).read() if rep.endswith(".csv") else subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    ki = hdr.index("Kernel Name")

    def val(r, metric):
        i = hdr.index(metric)
        return float(r[i].replace(",", "")) * SCALE.get(units[i], 1.0)

    acc = collections.defaultdict(list)
    for r in rows[2:]:
        name = next((n for n in NAMES if n in r[ki]), None)
        if name:
            acc[name].append(r)
    facts, traffic = {}, {}
    for name, rs in acc.items():
        facts[name] = {k: round(sum(val(r, m) for r in rs) / len(rs), 2) for k, m in FACTS.items()}
        facts[name]["launches_captured"] = len(rs)
        traffic[name] = int(sum(val(r, "dram__bytes_read.sum") + val(r, "dram__bytes_write.sum") for r in rs) / len(rs))
    for fname, new in (("ncu_kernel_facts.json", facts), ("ncu_traffic.json", traffic)):
        path = os.path.join(ROOT, "profiles", fname)
        cur = json.load(open(path)) if os.path.exists(path) else {}
        cur.setdefault(workload, {}).update(new)
        json.dump(cur, open(path, "w"), indent=1)
        print(fname, workload, json.dumps(new))


if __name__ == "__main__":
    main()
