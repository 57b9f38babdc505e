#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_linear.py -m gpu -q > gpurun_out/pytest_lin.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_lin.log
grep -E "passed|failed|Error|^E  " gpurun_out/pytest_lin.log | cut -c1-300 | tail -30
timeout 300 python scripts/lin_bench.py > gpurun_out/lin_bench.log 2>&1; cat gpurun_out/lin_bench.log
