#!/bin/bash
# Round 2 evidence: measured DRAM bytes + L2 hit rate of the products-shaped aggregation launches
# (F = 100 and F = 256, half graph on one GPU) and the tensor-pipe activity of vq_tc_kernel at 1M x {4,096; 16,384}
# x {128, 256, 512}.  Each ncu command runs only after the same command has exited 0 without the profiler.
mkdir -p gpurun_out
M="dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,gpu__time_duration.sum,sm__warps_active.avg.pct_of_peak_sustained_active"
timeout 600 python scripts/agg_bench.py > gpurun_out/agg_plain.log 2>&1 || { echo "agg_bench failed"; exit 1; }
timeout 1200 ncu --metrics $M --clock-control none --kernel-name-base demangled -k regex:'aggregate_(kernel|stream_kernel)' \
    --csv --log-file gpurun_out/ncu_agg_products.csv python scripts/agg_bench.py > gpurun_out/ncu_agg.log 2>&1
echo "ncu agg rc=$?"
T="sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__warps_active.avg.pct_of_peak_sustained_active"
: > gpurun_out/vq_points.log
for K in 4096 16384; do for D in 128 256 512; do
  timeout 300 python scripts/vq_stats.py 1000000 $K $D 3 >> gpurun_out/vq_points.log 2>&1 || { echo "vq_stats failed $K $D"; continue; }
  timeout 600 ncu --metrics $T --clock-control none --kernel-name-base demangled -k regex:vq_tc_kernel -c 2 \
      --csv --log-file gpurun_out/ncu_vq_${K}_${D}.csv python scripts/vq_stats.py 1000000 $K $D 1 > /dev/null 2>&1
done; done
cat gpurun_out/vq_points.log
grep -h "vq_tc_kernel" gpurun_out/ncu_vq_*.csv | cut -d, -f5,13- | cut -c1-200 | head -80
