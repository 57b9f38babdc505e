#!/bin/bash
# Round 2, call 2: the no-alloc entry point alone in a fresh process, the whole GPU suite (no -x: every failure
# shows), the aggregation A/B table, the launch list and the --set full capture of one eager bench step.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q --tb=short > gpurun_out/pytest_no_alloc_alone.log 2>&1
echo "rc=$?" >> gpurun_out/pytest_no_alloc_alone.log; tail -4 gpurun_out/pytest_no_alloc_alone.log
timeout 1500 python -m pytest tests -m gpu -q --tb=short > gpurun_out/pytest_gpu.log 2>&1
echo "rc=$?" >> gpurun_out/pytest_gpu.log; tail -8 gpurun_out/pytest_gpu.log
: > gpurun_out/agg_ab.log
for cfg in "STEMGNN_AGG_STREAM=0" "STEMGNN_AGG_STREAM=1 STEMGNN_AGG_HINTS=0" "STEMGNN_AGG_STREAM=1 STEMGNN_AGG_HINTS=1" \
           "STEMGNN_AGG_STREAM=128 STEMGNN_AGG_HINTS=1"; do
  echo "== $cfg" >> gpurun_out/agg_ab.log
  env $cfg timeout 300 python scripts/agg_bench.py >> gpurun_out/agg_ab.log 2>&1
done
grep -E "^==|fwd_us" gpurun_out/agg_ab.log | cut -c1-200
B="python bench.py --steps 1 --warmup 3 --no-cpu --no-graph --warm-seconds 0"
timeout 300 $B --dump-sequence gpurun_out/step_sequence.json > gpurun_out/bench_eager1.log 2>&1 || { echo "eager bench failed"; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k regex:stemgnn \
    -s 438 -c 160 --csv --log-file gpurun_out/launches_step.csv $B > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
timeout 1500 ncu --set full --clock-control none --kernel-name-base demangled -k regex:stemgnn -s 438 -c 160 \
    -o gpurun_out/prof_step -f $B > gpurun_out/ncu_step.log 2>&1
echo "ncu full rc=$?"
ncu -i gpurun_out/prof_step.ncu-rep --page raw --csv > gpurun_out/prof_step_raw.csv 2>/dev/null
ls -la gpurun_out/
[ $(stat -c %s gpurun_out/prof_step.ncu-rep) -gt 40000000 ] && rm -f gpurun_out/prof_step.ncu-rep
true
