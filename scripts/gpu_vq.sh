#!/bin/bash
# VQ search kernel: parity tests, then the (rows, K, D) timing points with per-kernel device times
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_vq_tc.py tests/test_gpu_modules.py -x -q -m gpu > gpurun_out/vq_tests.log 2>&1; echo "tests exit $?"; tail -3 gpurun_out/vq_tests.log
{
for d in 128 256 512; do for k in 1024 16384; do VQ_KERNELS=1 timeout 300 python scripts/vq_stats.py 1000000 $k $d 2 2>&1 | grep -v "Warn\|_warn_once"; done; done
VQ_KERNELS=1 timeout 300 python scripts/vq_stats.py 169343 512 256 3 2>&1 | grep -v "Warn\|_warn_once"
} > gpurun_out/vq_stats.log 2>&1; cat gpurun_out/vq_stats.log
