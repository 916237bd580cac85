#!/bin/bash
# Developer tool: A/B builds of the library (same ABI, different compile-time options) on the three workloads.
# usage: tools/ab_libs.sh default build_ab/libmoss_x.so ...     -> one line per (lib, workload): graph-replayed ms/step
for lib in "$@"; do
  for wl in "zju --steps 200 --warmup 20" "zju --views-per-step 23 --steps 40 --warmup 5" "stress --steps 30 --warmup 5" "monocap --steps 60 --warmup 5"; do
    set -- $wl
    if [ "$lib" = default ]; then unset MOSS_SPLAT_LIB; else export MOSS_SPLAT_LIB=$PWD/$lib; fi
    MOSS_PUBLIC_API=python MOSS_NO_SAMPLER=1 python bench.py --workload "$@" --no-extra --no-cpu-baseline --quick 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
k=d['roofline']['kernels_us']
print('$lib', '$wl', 'ms', d['ms_per_step'], 'bf', k.get('blend_forward'), 'bb', k.get('blend_backward'), 'k1', k.get('preprocess_forward'), 'k7', k.get('preprocess_backward'), 'sort', k.get('tile_sort'))
"
  done
done
