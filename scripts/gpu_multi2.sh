#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512"
timeout 200 $TR bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_arxiv_graph_$N.log 2>&1; echo "rc=$?" >> gpurun_out/bench_arxiv_graph_$N.log
grep -E "^\{|^rc=" gpurun_out/bench_arxiv_graph_$N.log | cut -c1-700
