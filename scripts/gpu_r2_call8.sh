#!/bin/bash
# Round 2: C1 / dense tests after the slab change, then compute-sanitizer (initcheck, racecheck, synccheck) over smoke()
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_configs.py tests/test_gpu_linear.py tests/test_gpu_modules.py -m gpu -q --tb=short > gpurun_out/pytest_dense.log 2>&1
echo "rc=$?" >> gpurun_out/pytest_dense.log; tail -6 gpurun_out/pytest_dense.log | cut -c1-250
for tool in initcheck racecheck synccheck; do
  timeout 900 bash scripts/gpu_r2_sanitizer.sh $tool > gpurun_out/sanitizer_${tool}_summary.log 2>&1
  tail -6 gpurun_out/sanitizer_${tool}_summary.log | cut -c1-250
done
