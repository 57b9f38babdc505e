#!/bin/bash
# full GPU suite + smoke + both bench arms + ncu launch list of the bench step
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
( time timeout 900 python bench.py --impl reference ) > gpurun_out/bench_ref.log 2>&1; grep -E "^\{|real" gpurun_out/bench_ref.log | cut -c1-400
( time timeout 900 python bench.py ) > gpurun_out/bench.log 2>&1; grep -E "^\{|real" gpurun_out/bench.log | cut -c1-600
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/bench_short.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv \
    --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu.log 2>&1
echo "ncu list rc=$?"
