#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513"
timeout 300 python -m pytest tests/test_gpu_partition.py tests/test_gpu_aggregate.py -m gpu -q -x > gpurun_out/pytest_part.log 2>&1; tail -2 gpurun_out/pytest_part.log
F=256 N=1200000 E=30000000 timeout 300 $TR scripts/check_partition_multi.py > gpurun_out/part_check_$N.log 2>&1; echo "rc=$?" >> gpurun_out/part_check_$N.log
grep -E "partition check|rc=|Error" gpurun_out/part_check_$N.log | tail -4
timeout 400 $TR bench.py --gpus $N --workload products --steps 5 --warmup 3 > gpurun_out/bench_products_$N.log 2>&1; echo "rc=$?" >> gpurun_out/bench_products_$N.log
grep -E "^\{|rc=" gpurun_out/bench_products_$N.log | cut -c1-330
