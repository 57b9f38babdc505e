#!/bin/bash
# N-GPU runs of BASELINE configs[2] (mix epoch, replicas over mini-batches) and configs[4] (VQ sweep,
# rows sharded): bash scripts/gpu_multi_c3c5.sh N [batches_per_rank]
N=${1:-2}
B=${2:-100}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
timeout 600 $TR scripts/mix_epoch.py --batches $B --cpu-batches 0 --trace 0 > gpurun_out/mix_epoch_${N}gpu.log 2>&1
grep '^{' gpurun_out/mix_epoch_${N}gpu.log | cut -c1-900
timeout 900 $TR scripts/vq_sweep_c5.py --cpu-rows 0 > gpurun_out/vq_sweep_c5_${N}gpu.log 2>&1
grep '^{' gpurun_out/vq_sweep_c5_${N}gpu.log | cut -c1-300
