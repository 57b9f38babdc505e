#!/bin/bash
# One compute-sanitizer tool per gpurun call (B200_PROFILING.md): bash scripts/gpu_r2_sanitizer.sh <initcheck|racecheck|synccheck|memcheck>
tool=${1:-initcheck}
mkdir -p gpurun_out
SMOKE='import __graft_entry__ as g; g.smoke()'
timeout 120 python -c "$SMOKE" > gpurun_out/sanitizer_plain_${tool}.log 2>&1 || { echo "plain smoke failed"; exit 1; }
timeout 1200 compute-sanitizer --tool $tool --print-limit 50 --kernel-name-exclude kernel_substring=at:: \
    python -c "$SMOKE" > gpurun_out/sanitizer_${tool}.log 2>&1
echo "sanitizer $tool rc=$?" >> gpurun_out/sanitizer_${tool}.log
grep -E "ERROR SUMMARY|smoke|rc=|Race|Uninit|hazard|Invalid" gpurun_out/sanitizer_${tool}.log | head -40
