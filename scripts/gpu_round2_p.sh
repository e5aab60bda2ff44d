#!/bin/bash
# round 2, GPU call P (N GPUs): multi-rank correctness script and the headline line after the NVLS broadcast / master gather change
set -x
N=${1:-2}
mkdir -p gpurun_out
timeout 300 python -m pytest tests/parallel/test_gpu_spmd.py -m gpu -x -q > gpurun_out/p${N}_spmd.log 2>&1; echo "spmd rc=$?"; tail -25 gpurun_out/p${N}_spmd.log
ELEPHAS_B200_TRACE=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 500)) bench.py --gpus $N --steps 20 --warmup 3 --sustained 0 > gpurun_out/p${N}_default.json 2> gpurun_out/p${N}_default.err
echo "rc=$?"; cut -c1-2500 gpurun_out/p${N}_default.json; grep -v "^>>>" gpurun_out/p${N}_default.err | tail -45 | cut -c1-200
