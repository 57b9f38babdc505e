#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_vq_tc.py tests/test_gpu_modules.py -m gpu -x -q > gpurun_out/pytest_tc.log 2>&1; echo "pytest_tc rc=$?" >> gpurun_out/pytest_tc.log
tail -8 gpurun_out/pytest_tc.log
timeout 600 python scripts/vq_stats.py > gpurun_out/vq_stats.log 2>&1; cat gpurun_out/vq_stats.log
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/bench.log 2>&1; echo "bench rc=$?" >> gpurun_out/bench.log
tail -2 gpurun_out/bench.log
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/bench_short.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'vq_tc_kernel' -s 3 -c 1 \
    -o gpurun_out/prof_tc2 -f python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
