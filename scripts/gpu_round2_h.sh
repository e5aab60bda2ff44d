#!/bin/bash
# round 2, GPU call H (N GPUs): multi-rank script (striped PS, per-batch CNN sync, NVLS), collective roofline, NVLS A/B in the step
set -x
N=${1:-2}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/parallel/test_gpu_spmd.py -m gpu -x -q > gpurun_out/h${N}_spmd.log 2>&1; echo "spmd rc=$?"; tail -25 gpurun_out/h${N}_spmd.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 500)) scripts/bench_collective.py > gpurun_out/h${N}_collective.jsonl 2> gpurun_out/h${N}_collective.err; echo "collective rc=$?"; cat gpurun_out/h${N}_collective.jsonl; tail -3 gpurun_out/h${N}_collective.err
run() { local name=$1; shift
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 500)) "$@" > gpurun_out/${name}.json 2> gpurun_out/${name}.err
  echo "$name rc=$?"; cut -c1-400 gpurun_out/${name}.json; tail -3 gpurun_out/${name}.err | cut -c1-300; }
run h${N}_default bench.py --gpus $N --steps 20 --warmup 3 --no-e2e --sustained 0
ELEPHAS_B200_NVLS=1 run h${N}_default_nvls bench.py --gpus $N --steps 20 --warmup 3 --no-e2e --sustained 0
run h${N}_mnist_async bench.py --gpus $N --config mnist_mlp --mode asynchronous --steps 2000 --warmup 20 --no-e2e
run h${N}_resnet bench.py --gpus $N --config resnet50
