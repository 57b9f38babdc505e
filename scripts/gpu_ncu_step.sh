#!/bin/bash
# one --set full capture of the main kernels of ONE bench step (eager launches), after the same
# command has run clean without the profiler
mkdir -p gpurun_out
timeout 600 python bench.py --steps 1 --warmup 3 --no-cpu --no-graph > gpurun_out/bench_short.log 2>&1 || exit 1
timeout 1200 ncu --set full --clock-control none --import-source on \
    --kernel-name-base demangled \
    -k regex:'aggregate_kernel<|lin_tc_kernel|wgrad_tc_kernel|bn::col_reduce_kernel|bn::apply_kernel|vq_tc_kernel|vq_gather_kernel' \
    -s ${NCU_SKIP:-48} -c ${NCU_COUNT:-24} -o gpurun_out/prof_step -f python bench.py --steps 1 --warmup 3 --no-cpu --no-graph > gpurun_out/ncu_step.log 2>&1
tail -2 gpurun_out/ncu_step.log
