#!/bin/bash
# round-2 call "s": loss kernel per-channel blocks A/B + tests + default bench
mkdir -p gpurun_out
O=gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > $O/s_gpu.txt 2>&1
timeout 600 python -m pytest tests/test_loss_gpu.py tests/test_zz_no_alloc.py tests/test_graphed_iteration_gpu.py -x -q -m gpu > $O/s_pytest_loss.log 2>&1
echo "pytest loss exit $?" >> $O/s_pytest_loss.log
for lib in default ab_libs/libmoss_s512_1.so ab_libs/libmoss_s1024_1.so ab_libs/libmoss_s256_2.so ab_libs/libmoss_old.so; do
  if [ "$lib" = default ]; then unset MOSS_SPLAT_LIB; else export MOSS_SPLAT_LIB=$PWD/$lib; fi
  echo "== $lib" >> $O/s_loss_ab.jsonl
  timeout 300 python tools/loss_bench.py 512x512 256x256 >> $O/s_loss_ab.jsonl 2>> $O/s_loss_ab.err
done
unset MOSS_SPLAT_LIB
timeout 300 python tools/iteration_bench.py > $O/s_iteration.jsonl 2> $O/s_iteration.err
timeout 900 python bench.py > $O/s_bench_default.json 2> $O/s_bench_default.err
echo "bench exit $?" >> $O/s_bench_default.err
timeout 900 python -m pytest tests -x -q -m gpu > $O/s_pytest_gpu.log 2>&1
echo "pytest exit $?" >> $O/s_pytest_gpu.log
tail -3 $O/s_pytest_loss.log; cat $O/s_loss_ab.jsonl | cut -c1-260; tail -3 $O/s_pytest_gpu.log
