#!/bin/bash
# Round 2, last call: suite (-x), smoke, light ncu pass over the aggregation launches of one eager step (the F <= 128
# kernel changed after the full capture), then both bench arms with the refreshed traffic file applied afterwards
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
grep -E "passed|failed|rc=" gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
B="python bench.py --steps 1 --warmup 3 --no-cpu --no-graph --no-partition --warm-seconds 0"
timeout 300 $B > /dev/null 2>&1 && timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum \
    --clock-control none --kernel-name-base demangled -k regex:aggregate_ -s 36 -c 6 --csv --log-file gpurun_out/ncu_aggregate_step.csv $B > /dev/null 2>&1
echo "ncu rc=$?"
( time timeout 900 python bench.py ) > gpurun_out/bench.log 2>&1; grep -E "^\{|real" gpurun_out/bench.log | cut -c1-200
