#!/bin/bash
# usage: gpu_r2_multi.sh NGPUS -- NCCL checks of the partitioned aggregation (bit-identical all-gather mode, owner-block
# overlapped modes within 1e-5) and the default bench line at N ranks with its nested partition record
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
N=${CHECK_N:-600000} E=${CHECK_E:-15000000} F=256 timeout 600 $TR scripts/check_partition_multi.py > gpurun_out/part_check_$N.log 2>&1
echo "rc=$?" >> gpurun_out/part_check_$N.log
grep -vE "^\[|NCCL|warn" gpurun_out/part_check_$N.log | tail -12
( time STEMGNN_HALO_BCAST=1 timeout 900 $TR bench.py --gpus $N --steps 100 --warmup 3 ) > gpurun_out/bench_arxiv_$N.log 2>&1
echo "rc=$?" >> gpurun_out/bench_arxiv_$N.log
grep -E "^\{|real|rc=" gpurun_out/bench_arxiv_$N.log | cut -c1-300
python - <<PY
import json
for l in open("gpurun_out/bench_arxiv_$N.log"):
    if l.startswith("{"):
        d = json.loads(l)
        p = d.get("partition") or {}
        print("ms/step", d["ms_per_step"], "e2e", d["e2e"]["ms_per_step"])
        print({k: v for k, v in p.items() if k != "modes"})
        for m in p.get("modes", []):
            print({k: (round(v, 3) if isinstance(v, float) else v) for k, v in m.items() if k != "note"})
PY
