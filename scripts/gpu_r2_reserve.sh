#!/bin/bash
# usage: gpu_r2_reserve.sh NGPUS -- the bench line's partition record with the dense kernels leaving 0 / 32 / 48 SMs
# to NCCL while an exchange is in flight
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
for r in 0 32 48; do
  STEMGNN_SM_RESERVE=$r timeout 600 $TR bench.py --gpus $N --steps 20 --warmup 3 --warm-seconds 0.2 > gpurun_out/bench_reserve${r}_$N.log 2>&1
  python - <<PY
import json
for l in open("gpurun_out/bench_reserve${r}_$N.log"):
    if l.startswith("{"):
        p = json.loads(l).get("partition") or {}
        print("reserve $r:", p.get("ms_per_step"), p.get("headline_mode"), p.get("error"))
        for m in p.get("modes", []):
            print("   ", {k: (round(v, 2) if isinstance(v, float) else v) for k, v in m.items() if k in ("mode", "ms_per_step", "halo_wait_ms", "aggregate_ms", "dense_ms")})
PY
done
