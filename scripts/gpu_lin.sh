#!/bin/bash
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'wgrad_tc_kernel|wgrad_reduce' -s 8 -c 2 -o gpurun_out/prof_wg -f python scripts/lin_bench.py > gpurun_out/ncu_lin.log 2>&1; tail -3 gpurun_out/ncu_lin.log
