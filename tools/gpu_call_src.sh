#!/bin/bash
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
N="ncu --set full --clock-control none --import-source on -f"
for k in blend_backward_kernel; do
  timeout 300 $N -k regex:$k -s 1 -c 1 -o /tmp/$k python tools/tiny_step.py stress 2 > gpurun_out/src_ncu_$k.log 2>&1
  ncu -i /tmp/$k.ncu-rep --page source --csv --print-source sass > gpurun_out/src_$k.csv 2>/dev/null
  ncu -i /tmp/$k.ncu-rep --page raw --csv > gpurun_out/src_$k.raw.csv 2>/dev/null
done
ls -la gpurun_out/src_blend_backward*
