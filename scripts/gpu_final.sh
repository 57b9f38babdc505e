#!/bin/bash
# what the driver runs at round end: GPU tests, smoke, both bench arms
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
( time timeout 900 python bench.py --impl reference ) > gpurun_out/bench_ref.log 2>&1; grep -E "^\{|real" gpurun_out/bench_ref.log | cut -c1-600
( time timeout 900 python bench.py ) > gpurun_out/bench.log 2>&1; grep -E "^\{|real" gpurun_out/bench.log | cut -c1-400
