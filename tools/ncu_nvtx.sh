#!/bin/bash
# NVTX ranges seen by a tool: the launch list of ONLY the backward call's kernels, selected by range name.
#   tools/ncu_nvtx.sh [scene] -> gpurun_out/nvtx_backward_launches.csv (+ the unfiltered list beside it)
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
scene=${1:-zju}
timeout 120 python tools/tiny_step.py $scene 2 > /dev/null 2>&1 || { echo PLAIN_FAILED; exit 1; }
MOSS_NVTX=1 timeout 300 ncu --nvtx --nvtx-include "moss_rasterize_backward/" --metrics gpu__time_duration.sum --clock-control none \
  --csv --log-file gpurun_out/nvtx_backward_launches.csv python tools/tiny_step.py $scene 2 > gpurun_out/nvtx_backward.log 2>&1
MOSS_NVTX=1 timeout 300 ncu --nvtx --metrics gpu__time_duration.sum --clock-control none --print-nvtx-rename kernel \
  --csv --log-file gpurun_out/nvtx_all_launches.csv python tools/tiny_step.py $scene 2 > gpurun_out/nvtx_all.log 2>&1
grep -c "moss::" gpurun_out/nvtx_backward_launches.csv gpurun_out/nvtx_all_launches.csv
cut -d, -f5 gpurun_out/nvtx_backward_launches.csv | sort | uniq -c
