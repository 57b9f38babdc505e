#!/bin/bash
# B-ring depth / A double-buffer sweep of the tcgen05 VQ search
mkdir -p gpurun_out
{
for d in 128 256; do
 for k in 512 4096 16384; do
  rows=1000000
  echo "== default D=$d K=$k"; python scripts/vq_stats.py $rows $k $d
  for st in 2 3; do echo "== stages<=$st"; STEMGNN_VQ_STAGES=$st python scripts/vq_stats.py $rows $k $d; done
  echo "== double A off"; STEMGNN_VQ_DOUBLE_A=0 python scripts/vq_stats.py $rows $k $d
  echo "== double A on"; STEMGNN_VQ_DOUBLE_A=1 python scripts/vq_stats.py $rows $k $d
 done
done
python scripts/vq_stats.py 1000000 16384 512
python scripts/vq_stats.py 1000000 16384 384
} > gpurun_out/vq_sweep.log 2>&1
cat gpurun_out/vq_sweep.log
