#!/bin/bash
# ncu --set full of the tcgen05 VQ search kernel at two widths (after a clean run of the same command)
mkdir -p gpurun_out
timeout 300 python scripts/vq_one.py 1000000 16384 256 1 || exit 1
timeout 300 python scripts/vq_one.py 1000000 16384 128 1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:vq_tc_kernel -c 1 -o gpurun_out/prof_vq256 -f python scripts/vq_one.py 1000000 16384 256 1 > gpurun_out/ncu_vq256.log 2>&1
tail -2 gpurun_out/ncu_vq256.log
