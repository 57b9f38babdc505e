#!/bin/bash
# Launch list of the bench step (our kernels + the rest) and one --set full capture of the top kernels.
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_vq_tc.py -m gpu -x -q > gpurun_out/pytest_tc.log 2>&1; echo "pytest_tc rc=$?" >> gpurun_out/pytest_tc.log
tail -5 gpurun_out/pytest_tc.log
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/bench_short.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv \
    --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu.log 2>&1
echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'vq_tc_kernel|aggregate_kernel|vq_gather_kernel' -s 12 -c 6 \
    -o gpurun_out/prof_r01 -f python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
