#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
for lib in ab_libs/libmoss_old.so default; do
  if [ "$lib" = default ]; then unset MOSS_SPLAT_LIB; else export MOSS_SPLAT_LIB=$PWD/$lib; fi
  echo "== $lib" >> $O/w_loss.jsonl
  timeout 300 python tools/loss_bench.py 640x480 1024x1024 1920x1080 >> $O/w_loss.jsonl 2>> $O/w_loss.err
done
cut -c1-200 $O/w_loss.jsonl
