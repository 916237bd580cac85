#!/bin/bash
# `ncu --set full` captures of the rasterizer's kernels at ZJU (one view, 23-view batch) and Stress size, exported as raw CSV
# (gpurun_out/<prefix>_ncu_full_{zju,zju_views23,stress}.raw.csv); fold them into profiles/ with tools/ncu_facts.py.
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
T=${1:-ncu}
N="ncu --set full --clock-control none --import-source on -f"
timeout 120 python tools/tiny_step.py zju 3 > /dev/null 2>&1 && timeout 200 python tools/tiny_step.py stress 2 > /dev/null 2>&1 || echo PLAIN_FAILED
timeout 400 $N -k regex:'preprocess|tile_scan|scatter_keys|tile_sort|blend_' -s 7 -c 14 -o /tmp/zju python tools/tiny_step.py zju 3 > gpurun_out/${T}_ncu_zju.log 2>&1
timeout 400 $N -k regex:'preprocess|tile_scan|scatter_keys|tile_sort|blend_' -s 7 -c 7 -o /tmp/stress python tools/tiny_step.py stress 2 > gpurun_out/${T}_ncu_stress.log 2>&1
timeout 400 $N -k regex:'preprocess|tile_scan|scatter_keys|tile_sort|blend_|sum_views' -s 8 -c 8 -o /tmp/v23 python tools/batch_step.py 23 2 > gpurun_out/${T}_ncu_v23.log 2>&1
ncu -i /tmp/zju.ncu-rep --page raw --csv > gpurun_out/${T}_ncu_full_zju.raw.csv 2>/dev/null
ncu -i /tmp/stress.ncu-rep --page raw --csv > gpurun_out/${T}_ncu_full_stress.raw.csv 2>/dev/null
ncu -i /tmp/v23.ncu-rep --page raw --csv > gpurun_out/${T}_ncu_full_zju_views23.raw.csv 2>/dev/null
wc -l gpurun_out/${T}_ncu_full_*.raw.csv
