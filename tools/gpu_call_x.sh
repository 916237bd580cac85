#!/bin/bash
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
O=gpurun_out
for lib in default ab_libs/libmoss_ns0.so ab_libs/libmoss_ns1.so; do
  if [ "$lib" = default ]; then unset MOSS_SPLAT_LIB; else export MOSS_SPLAT_LIB=$PWD/$lib; fi
  echo "== $lib" >> $O/x_loss.jsonl
  timeout 300 python tools/loss_bench.py 640x480 1024x1024 1920x1080 >> $O/x_loss.jsonl 2>> $O/x_loss.err
done
cut -c1-200 $O/x_loss.jsonl
export MOSS_SPLAT_LIB=$PWD/ab_libs/libmoss_ns1.so
timeout 300 ncu --set full --clock-control none -f -k regex:image_losses -c 1 -o /tmp/l1 python tools/loss_step.py 1920x1080 1 > $O/x_ncu.log 2>&1
ncu -i /tmp/l1.ncu-rep --page details 2>/dev/null | grep -E "Block Limit|Occupancy|Duration|Shared Memory Config|Registers Per" > $O/x_ncu_ns1.txt
export MOSS_SPLAT_LIB=$PWD/ab_libs/libmoss_old.so
timeout 300 ncu --set full --clock-control none -f -k regex:image_losses -c 1 -o /tmp/l2 python tools/loss_step.py 1920x1080 1 >> $O/x_ncu.log 2>&1
ncu -i /tmp/l2.ncu-rep --page details 2>/dev/null | grep -E "Block Limit|Occupancy|Duration|Shared Memory Config|Registers Per" > $O/x_ncu_old.txt
cat $O/x_ncu_ns1.txt $O/x_ncu_old.txt
