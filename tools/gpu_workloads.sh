#!/bin/bash
# MonoCap and Stress lines of both arms on one box (BASELINE.json configs 3 and 5), plus the 23-view batch and the
# forward-only Stress render: gpurun_out/<prefix>_bench_{ours,reference}_{monocap,stress,...}.json
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
T=${1:-wl}
O=gpurun_out
for wl in monocap stress; do
  steps=60; [ $wl = stress ] && steps=30
  timeout 600 python bench.py --impl reference --workload $wl --steps $steps --warmup 5 > $O/${T}_bench_reference_$wl.json 2> $O/${T}_bench_reference_$wl.err
  timeout 600 python bench.py --workload $wl --steps $steps --warmup 5 --no-extra --no-cpu-baseline > $O/${T}_bench_ours_$wl.json 2> $O/${T}_bench_ours_$wl.err
done
timeout 600 python bench.py --workload stress --forward-only --steps 30 --warmup 5 --no-extra --no-cpu-baseline > $O/${T}_bench_ours_stress_fwd.json 2> $O/${T}_bench_ours_stress_fwd.err
timeout 600 python bench.py --impl reference --workload stress --forward-only --steps 30 --warmup 5 > $O/${T}_bench_reference_stress_fwd.json 2> $O/${T}_bench_reference_stress_fwd.err
timeout 600 python bench.py --views-per-step 23 --steps 40 --warmup 5 --no-extra --no-cpu-baseline > $O/${T}_bench_ours_zju_views23.json 2> $O/${T}_bench_ours_zju_views23.err
python - <<EOF
import json, glob
for f in sorted(glob.glob("$O/${T}_bench_*.json")):
    try:
        d = json.loads([l for l in open(f).read().splitlines() if l.startswith("{")][-1])
        print(f.split("/")[-1], d.get("metric"), d.get("value"), d.get("ms_per_step"), (d.get("e2e") or {}).get("ms_per_step"))
    except Exception as e:
        print(f, "unreadable", e)
EOF
