#!/bin/bash
# usage: gpu_multi.sh NGPUS
N=${1:-2}
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR scripts/check_partition_multi.py > gpurun_out/part_check_$N.log 2>&1; echo "rc=$?" >> gpurun_out/part_check_$N.log
tail -3 gpurun_out/part_check_$N.log
timeout 900 $TR bench.py --gpus $N --workload products --steps 5 --warmup 3 > gpurun_out/bench_products_$N.log 2>&1; echo "rc=$?" >> gpurun_out/bench_products_$N.log
tail -2 gpurun_out/bench_products_$N.log
timeout 600 $TR bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_arxiv_$N.log 2>&1; echo "rc=$?" >> gpurun_out/bench_arxiv_$N.log
tail -2 gpurun_out/bench_arxiv_$N.log
