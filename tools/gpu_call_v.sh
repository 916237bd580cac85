#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python -m pytest tests/test_loss_gpu.py tests/test_zz_no_alloc.py tests/test_graphed_iteration_gpu.py -x -q -m gpu > $O/v_pytest_loss.log 2>&1
echo "pytest loss exit $?" >> $O/v_pytest_loss.log
timeout 300 python tools/loss_bench.py 512x512 256x256 640x480 1024x1024 1920x1080 > $O/v_loss.jsonl 2> $O/v_loss.err
tail -4 $O/v_pytest_loss.log; cut -c1-230 $O/v_loss.jsonl
