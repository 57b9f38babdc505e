#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_bn_act.py tests/test_gpu_linear.py -m gpu -q > gpurun_out/pytest_bn.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_bn.log
grep -E "passed|failed|Error|^E  " gpurun_out/pytest_bn.log | cut -c1-300 | tail -30
timeout 900 python bench.py --steps 20 --warmup 5 --no-cpu > gpurun_out/bench_bn.log 2>&1; tail -1 gpurun_out/bench_bn.log | cut -c1-300
