#!/bin/bash
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
N="ncu --set full --clock-control none --import-source on -f"
timeout 120 python tools/loss_step.py 512x512 3 > gpurun_out/u_plain.log 2>&1 || echo PLAIN_FAILED
timeout 300 $N -k regex:image_losses -c 2 -o /tmp/loss python tools/loss_step.py 512x512 2 > gpurun_out/u_ncu.log 2>&1
ncu -i /tmp/loss.ncu-rep --page raw --csv > gpurun_out/u_loss.raw.csv 2>/dev/null
ncu -i /tmp/loss.ncu-rep --page source --csv --print-source sass > gpurun_out/u_loss.source.csv 2>/dev/null
ncu -i /tmp/loss.ncu-rep --page details > gpurun_out/u_loss.details.txt 2>/dev/null
ls -la gpurun_out | tail -8
