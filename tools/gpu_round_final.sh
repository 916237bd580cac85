#!/bin/bash
# Round-end evidence on one B200 box: full GPU test suite, smoke(), both bench arms back to back, the ncu launch list of the
# same bench command and one --set full capture; everything lands in gpurun_out/ with the prefix given as $1.
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
T=${1:-final}
O=gpurun_out
timeout 1200 python -m pytest tests -x -q -m gpu > $O/${T}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/${T}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/${T}_smoke.log 2>&1; echo "smoke exit $?" >> $O/${T}_smoke.log
timeout 900 python bench.py --impl reference > $O/${T}_bench_reference_zju.json 2> $O/${T}_bench_reference.err; echo "ref exit $?" >> $O/${T}_bench_reference.err
timeout 900 python bench.py > $O/${T}_bench_ours_zju_default.json 2> $O/${T}_bench_ours.err; echo "ours exit $?" >> $O/${T}_bench_ours.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${T}_launches_zju.csv \
  python bench.py --quick --no-graph --no-extra --no-cpu-baseline > $O/${T}_ncu_launches.log 2>&1
tail -2 $O/${T}_pytest_gpu.log; tail -2 $O/${T}_smoke.log; tail -1 $O/${T}_bench_reference.err; tail -1 $O/${T}_bench_ours.err
python - <<EOF
import json
for f in ("$O/${T}_bench_reference_zju.json", "$O/${T}_bench_ours_zju_default.json"):
    try:
        d = json.loads([l for l in open(f).read().splitlines() if l.startswith("{")][-1])
        print(f, d.get("value"), d.get("ms_per_step"), (d.get("e2e") or {}).get("ms_per_step"), d.get("roofline", {}).get("frac"))
    except Exception as e:
        print(f, "unreadable", e)
EOF
