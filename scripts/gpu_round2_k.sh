#!/bin/bash
# round 2, GPU call K (N GPUs): the headline line with the phase tracer on (where does an end-to-end fit spend its wall clock?)
set -x
N=${1:-8}
mkdir -p gpurun_out
ELEPHAS_B200_TRACE=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 500)) bench.py --gpus $N --steps 20 --warmup 3 --sustained 0 > gpurun_out/k${N}_default.json 2> gpurun_out/k${N}_default.err
echo "rc=$?"; cut -c1-2500 gpurun_out/k${N}_default.json; grep -v "^>>>" gpurun_out/k${N}_default.err | tail -60 | cut -c1-200
