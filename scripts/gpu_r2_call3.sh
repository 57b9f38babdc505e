#!/bin/bash
# Round 2, call 3: the whole GPU suite on the new build (K0 bucket build, sparse adjacency, owner-block exchange),
# smoke, the default bench line (nested partition record at N = 1 included) and the reference arm.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --tb=short > gpurun_out/pytest_gpu.log 2>&1
echo "rc=$?" >> gpurun_out/pytest_gpu.log; tail -25 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log; tail -2 gpurun_out/smoke.log
( time timeout 900 python bench.py ) > gpurun_out/bench.log 2>&1; echo "bench rc=$?" >> gpurun_out/bench.log
grep -E "^\{|real|rc=" gpurun_out/bench.log | cut -c1-1500
( time timeout 600 python bench.py --impl reference --steps 5 --warmup 1 ) > gpurun_out/bench_ref.log 2>&1
grep -E "^\{|real" gpurun_out/bench_ref.log | cut -c1-600
