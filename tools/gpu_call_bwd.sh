#!/bin/bash
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_batch_gpu.py tests/test_fused_gpu.py tests/test_zy_fuzz_gpu.py -x -q -m gpu > gpurun_out/bwd_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/bwd_pytest.log
tail -3 gpurun_out/bwd_pytest.log
bash tools/ab_libs.sh default > gpurun_out/bwd_ab.txt 2> gpurun_out/bwd_ab.err
cat gpurun_out/bwd_ab.txt
