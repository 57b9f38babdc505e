#!/bin/bash
# Round 2, final single-GPU call: GPU suite (as the driver runs it: -x), smoke, both bench arms, then the launch list
# and the --set full capture of one eager step of the FINAL kernels (-> scripts/ncu_traffic.py -> profiles/traffic.json)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
grep -E "passed|failed|^FAILED|rc=" gpurun_out/pytest_gpu.log | head
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log; tail -2 gpurun_out/smoke.log
( time timeout 900 python bench.py --impl reference ) > gpurun_out/bench_ref.log 2>&1; grep -E "^\{|real" gpurun_out/bench_ref.log | cut -c1-300
( time timeout 900 python bench.py ) > gpurun_out/bench.log 2>&1; grep -E "^\{|real" gpurun_out/bench.log | cut -c1-300
B="python bench.py --steps 1 --warmup 3 --no-cpu --no-graph --no-partition --warm-seconds 0"
timeout 300 $B --dump-sequence gpurun_out/step_sequence.json > gpurun_out/bench_eager1.log 2>&1 || { echo "eager bench failed"; exit 1; }
timeout 1500 ncu --set full --clock-control none --kernel-name-base demangled -k regex:stemgnn -s 354 -c 130 \
    -o gpurun_out/prof_step -f $B > gpurun_out/ncu_step.log 2>&1
echo "ncu full rc=$?"
ncu -i gpurun_out/prof_step.ncu-rep --page raw --csv > gpurun_out/prof_step_raw.csv 2>/dev/null
rm -f gpurun_out/prof_step.ncu-rep
ls -la gpurun_out/prof_step_raw.csv gpurun_out/step_sequence.json
