#!/bin/bash
# N-GPU run: products-shaped partitioned step with the halo-exchange A/B, then the default (replica) bench
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
STEMGNN_HALO_SWEEP="2:f32,4:f32,1:bf16,2:bf16" timeout 600 $TR bench.py --gpus $N --workload products --steps 5 --warmup 3 > gpurun_out/bench_products_${N}_sweep.log 2>&1; echo "rc=$?" >> gpurun_out/bench_products_${N}_sweep.log
grep -E "^\{|rc=|sweep" gpurun_out/bench_products_${N}_sweep.log | cut -c1-330
timeout 300 $TR bench.py --gpus $N > gpurun_out/bench_arxiv_${N}.log 2>&1; echo "rc=$?" >> gpurun_out/bench_arxiv_${N}.log
grep -E "^\{|rc=" gpurun_out/bench_arxiv_${N}.log | cut -c1-330
