#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_partition.py -m gpu -x -q > gpurun_out/pytest_part.log 2>&1; echo "pytest_part rc=$?" >> gpurun_out/pytest_part.log
tail -8 gpurun_out/pytest_part.log
timeout 900 python bench.py --workload products --steps 5 --warmup 3 > gpurun_out/bench_products_1.log 2>&1; echo "rc=$?" >> gpurun_out/bench_products_1.log
tail -3 gpurun_out/bench_products_1.log
