#!/bin/bash
# K9 / K10: new GPU tests first, then the whole GPU suite, then the stand-alone timings
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_augment.py tests/test_gpu_sampler.py -q -x > gpurun_out/pytest_aug.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_aug.log
tail -25 gpurun_out/pytest_aug.log
for w in arxiv mix products; do timeout 300 python scripts/aug_bench.py $w; done > gpurun_out/aug_bench.log 2>&1
cat gpurun_out/aug_bench.log | tail -30
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
