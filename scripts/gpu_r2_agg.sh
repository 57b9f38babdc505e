#!/bin/bash
# Round 2: the staged-index aggregation kernel -- parity tests, then A/B timings (scripts/agg_bench.py)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_aggregate.py tests/test_gpu_partition.py tests/test_gpu_c2_parity.py \
    tests/test_gpu_reference_wrappers.py tests/test_gpu_fullsize.py -m gpu -q --tb=short > gpurun_out/pytest_agg.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_agg.log
tail -30 gpurun_out/pytest_agg.log
: > gpurun_out/agg_ab.log
for cfg in "STEMGNN_AGG_STREAM=0" "STEMGNN_AGG_STREAM=1 STEMGNN_AGG_HINTS=0" "STEMGNN_AGG_STREAM=1 STEMGNN_AGG_HINTS=1" \
           "STEMGNN_AGG_STREAM=1 STEMGNN_AGG_HINTS=1 STEMGNN_AGG_STREAM_OCC=4"; do
  echo "== $cfg" >> gpurun_out/agg_ab.log
  env $cfg timeout 600 python scripts/agg_bench.py ${AGG_ARGS:-} >> gpurun_out/agg_ab.log 2>&1
done
cat gpurun_out/agg_ab.log
