#!/bin/bash
# tcgen05 VQ path bring-up: new parity tests first (bounded by timeout), then the rest.
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_vq_tc.py -m gpu -x -q > gpurun_out/pytest_tc.log 2>&1; echo "pytest_tc rc=$?" >> gpurun_out/pytest_tc.log
tail -30 gpurun_out/pytest_tc.log
timeout 900 python -m pytest tests -m gpu -q --deselect tests/test_gpu_vq_tc.py > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
tail -3 gpurun_out/smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/bench.log 2>&1; echo "bench rc=$?" >> gpurun_out/bench.log
tail -3 gpurun_out/bench.log
