#!/bin/bash
# Round 2: suite with the fused SAGE layer (ABI 6), bench A/B fused vs composed, then the evidence captures
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --tb=short > gpurun_out/pytest_gpu.log 2>&1
echo "rc=$?" >> gpurun_out/pytest_gpu.log; grep -E "passed|failed|^FAILED|^E  " gpurun_out/pytest_gpu.log | head -20
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log; tail -2 gpurun_out/smoke.log
STEMGNN_SAGE_FUSED=0 timeout 600 python bench.py --no-cpu --no-partition > gpurun_out/bench_composed.log 2>&1
timeout 900 python bench.py > gpurun_out/bench.log 2>&1; echo "bench rc=$?" >> gpurun_out/bench.log
python - <<'PY'
import json
for f in ("gpurun_out/bench_composed.log", "gpurun_out/bench.log"):
    for l in open(f):
        if l.startswith("{"):
            d = json.loads(l)
            print(f, "ms/step", round(d["ms_per_step"], 4), "e2e", round(d["e2e"]["ms_per_step"], 4), "launches", d["gpu_launches"])
            print("   ", {k: round(v["ms_per_step"], 4) for k, v in d["kernels"].items()})
            if d.get("partition"): print("    partition", d["partition"].get("ms_per_step"), d["partition"].get("error"))
PY
bash scripts/gpu_r2_evidence.sh > gpurun_out/evidence.log 2>&1; tail -40 gpurun_out/evidence.log | cut -c1-220
