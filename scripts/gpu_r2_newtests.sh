#!/bin/bash
# Round 2: the new parity tests (C2 whole step, full-shape fixtures, reference wrappers, bit stability)
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_c2_parity.py tests/test_gpu_vq_stability.py tests/test_gpu_reference_wrappers.py \
    -m gpu -q -rs --tb=short ${PYTEST_ARGS:-} > gpurun_out/pytest_newtests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_newtests.log
tail -60 gpurun_out/pytest_newtests.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
tail -3 gpurun_out/smoke.log
