#!/bin/bash
# Round 2: suite after the blocks-mode fix, bench A/B of the reversed sweeps (STEMGNN_L2_PINGPONG), the launch list of
# one replayed step (K0 kernels), and the K = 16,384 x D = 256 VQ point again
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --tb=short > gpurun_out/pytest_gpu.log 2>&1
echo "rc=$?" >> gpurun_out/pytest_gpu.log; grep -E "passed|failed|^FAILED|^E  " gpurun_out/pytest_gpu.log | head -20
for pp in 0 1; do
  STEMGNN_L2_PINGPONG=$pp timeout 600 python bench.py --no-cpu --no-partition > gpurun_out/bench_pp$pp.log 2>&1
done
python - <<'PY'
import json
for f in ("gpurun_out/bench_pp0.log", "gpurun_out/bench_pp1.log"):
    for l in open(f):
        if l.startswith("{"):
            d = json.loads(l)
            print(f, "ms/step", round(d["ms_per_step"], 4), "e2e", round(d["e2e"]["ms_per_step"], 4), "launches", d["gpu_launches"])
            print("   ", {k: round(v["ms_per_step"], 4) for k, v in d["kernels"].items()})
PY
B="python bench.py --steps 2 --warmup 3 --no-cpu --no-partition --warm-seconds 0"
timeout 300 $B > gpurun_out/bench_short.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none \
    --kernel-name-base demangled -k regex:stemgnn -s 300 -c 200 --csv --log-file gpurun_out/launches_step_final.csv $B > gpurun_out/ncu_launches.log 2>&1
grep -E "k0_|radix" gpurun_out/launches_step_final.csv | cut -d, -f5,9,15 | head -12
for i in 1 2; do timeout 300 python scripts/vq_stats.py 1000000 16384 256 5; done
VQ_KERNELS=1 timeout 300 python scripts/vq_stats.py 1000000 16384 256 3
