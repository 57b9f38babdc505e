#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_linear.py tests/test_gpu_bn_act.py -m gpu -q > gpurun_out/pytest_lin.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_lin.log
grep -E "passed|failed|Error|^E  " gpurun_out/pytest_lin.log | cut -c1-300 | tail -10
STEMGNN_LIN_BM=256 timeout 600 python -m pytest tests/test_gpu_linear.py -m gpu -q -k "linear" > gpurun_out/pytest_lin256.log 2>&1; tail -1 gpurun_out/pytest_lin256.log
timeout 300 python scripts/lin_bench.py > gpurun_out/lin_bench128.log 2>&1; head -6 gpurun_out/lin_bench128.log
STEMGNN_LIN_BM=256 timeout 300 python scripts/lin_bench.py > gpurun_out/lin_bench256.log 2>&1; head -6 gpurun_out/lin_bench256.log
