#!/bin/bash
# usage: gpu_r2_prio.sh NGPUS -- the partitioned aggregation check with the NCCL stream at default / high priority
N=${1:-4}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
for prio in 0 1; do
  STEMGNN_NCCL_HIGH_PRIO=$prio N=1200000 E=30000000 F=256 timeout 600 $TR scripts/check_partition_multi.py > gpurun_out/part_prio${prio}_$N.log 2>&1
  echo "== high_prio=$prio rc=$?"; grep -E "partition check|blocks/" gpurun_out/part_prio${prio}_$N.log
done
NCCL_DEBUG=INFO N=1200000 E=30000000 F=256 timeout 600 $TR scripts/check_partition_multi.py 2>&1 | grep -iE "NVLS|algo|channels|Connected|via" | sort | uniq -c | sort -rn | head -20
