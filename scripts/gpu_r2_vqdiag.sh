#!/bin/bash
# Round 2, call 1: is the VQ assignment bit-stable run to run (poisoned workspaces), and does smoke()
# reproduce the driver's `vq indices mismatch` plain / under ncu?  Logs -> gpurun_out/.
mkdir -p gpurun_out
SMOKE='import __graft_entry__ as g; g.smoke()'
timeout 900 python scripts/vq_stability.py 50 > gpurun_out/vq_stability.log 2>&1; echo "rc=$?" >> gpurun_out/vq_stability.log
tail -12 gpurun_out/vq_stability.log
: > gpurun_out/smoke_plain.log
for i in $(seq 1 15); do
  timeout 120 python -c "$SMOKE" >> gpurun_out/smoke_plain.log 2>&1; echo "run $i rc=$?" >> gpurun_out/smoke_plain.log
done
grep -c "smoke ok" gpurun_out/smoke_plain.log
: > gpurun_out/smoke_ncu.log
for i in 1 2 3 4; do
  timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 1000 --csv \
      --log-file gpurun_out/smoke_ncu_launches_$i.csv python -c "$SMOKE" >> gpurun_out/smoke_ncu.log 2>&1
  echo "ncu run $i rc=$?" >> gpurun_out/smoke_ncu.log
done
for i in 5 6; do
  timeout 300 ncu --metrics gpu__time_duration.sum -c 1000 --csv \
      --log-file gpurun_out/smoke_ncu_launches_$i.csv python -c "$SMOKE" >> gpurun_out/smoke_ncu.log 2>&1
  echo "ncu (default clock control) run $i rc=$?" >> gpurun_out/smoke_ncu.log
done
grep -E "smoke|rc=" gpurun_out/smoke_ncu.log | tail -30
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
