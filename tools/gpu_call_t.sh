#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
for lib in default ab_libs/libmoss_s512_1.so; do
  if [ "$lib" = default ]; then unset MOSS_SPLAT_LIB; else export MOSS_SPLAT_LIB=$PWD/$lib; fi
  echo "== $lib" >> $O/t_loss_ab.jsonl
  timeout 300 python tools/loss_bench.py 512x512 256x256 640x480 >> $O/t_loss_ab.jsonl 2>> $O/t_loss_ab.err
done
cut -c1-230 $O/t_loss_ab.jsonl
