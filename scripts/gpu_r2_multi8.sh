#!/bin/bash
# usage: gpu_r2_multi8.sh NGPUS -- partition check (all modes, high-priority NCCL stream) at products-like density,
# an NCCL_ALGO sweep of the blocking all-gather mode, then the default bench line with its nested partition record
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
STEMGNN_NCCL_HIGH_PRIO=1 N=2449029 E=61859140 F=256 timeout 900 $TR scripts/check_partition_multi.py > gpurun_out/part_check_products_$N.log 2>&1
echo "rc=$?" >> gpurun_out/part_check_products_$N.log
grep -E "partition check|blocks/|chunks/|rc=" gpurun_out/part_check_products_$N.log
for algo in Ring NVLS; do
  NCCL_ALGO=$algo STEMGNN_CHECK_GATHER_ONLY=1 N=2449029 E=61859140 F=256 timeout 600 $TR scripts/check_partition_multi.py 2>&1 | grep -E "partition check" | sed "s/^/NCCL_ALGO=$algo /"
done
( time timeout 900 $TR bench.py --gpus $N --steps 100 --warmup 3 ) > gpurun_out/bench_arxiv_$N.log 2>&1
echo "rc=$?" >> gpurun_out/bench_arxiv_$N.log
python - <<PY
import json
for l in open("gpurun_out/bench_arxiv_$N.log"):
    if l.startswith("{"):
        d = json.loads(l)
        p = d.get("partition") or {}
        print("ms/step", d["ms_per_step"], "e2e", d["e2e"]["ms_per_step"])
        print({k: v for k, v in p.items() if k not in ("modes", "parity", "workload", "l2")})
        for m in p.get("modes", []):
            print({k: (round(v, 3) if isinstance(v, float) else v) for k, v in m.items() if k not in ("note", "halo_mode")})
PY
