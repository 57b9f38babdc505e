#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_modules.py -m gpu -q -x > gpurun_out/pytest_mod.log 2>&1; tail -2 gpurun_out/pytest_mod.log
timeout 900 python bench.py --steps 20 --warmup 5 --no-cpu > gpurun_out/bench_graph.log 2>&1; tail -1 gpurun_out/bench_graph.log | cut -c1-1800
timeout 900 python bench.py --steps 20 --warmup 5 --no-cpu --no-graph > gpurun_out/bench_eager.log 2>&1; tail -1 gpurun_out/bench_eager.log | cut -c1-700
