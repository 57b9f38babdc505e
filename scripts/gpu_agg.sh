#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_aggregate.py tests/test_gpu_partition.py tests/test_gpu_modules.py -m gpu -x -q > gpurun_out/pytest_agg.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_agg.log
tail -5 gpurun_out/pytest_agg.log
timeout 600 python scripts/agg_bench.py fast > gpurun_out/agg_fast.log 2>&1; tail -8 gpurun_out/agg_fast.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/bench_arxiv_fast.log 2>&1; tail -1 gpurun_out/bench_arxiv_fast.log | cut -c1-400
cp stem-gnn_b200/libstemgnn_b200.so /tmp/lib_default.so
for mb in occ; do
  cp build/alt/lib_$mb.so stem-gnn_b200/libstemgnn_b200.so
  timeout 600 python scripts/agg_bench.py mb$mb > gpurun_out/agg_mb$mb.log 2>&1; tail -8 gpurun_out/agg_mb$mb.log
done
cp /tmp/lib_default.so stem-gnn_b200/libstemgnn_b200.so
