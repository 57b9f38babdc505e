#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_linear.py -m gpu -q -k "large" > gpurun_out/pytest_lin.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_lin.log
grep -E "passed|failed|Error|^E  " gpurun_out/pytest_lin.log | cut -c1-300 | tail -20
timeout 900 python bench.py --workload products --steps 5 --warmup 3 --no-cpu > gpurun_out/bench_products_1.log 2>&1; echo "rc=$?" >> gpurun_out/bench_products_1.log
tail -2 gpurun_out/bench_products_1.log | cut -c1-600
