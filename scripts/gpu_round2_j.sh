#!/bin/bash
# round 2, GPU call J (N GPUs): the driver's headline line at N, NVLS A/B, collective roofline, host-link ceiling, parameter-server
# scaling, the multi-rank correctness script, ResNet-50 per-batch
set -x
N=${1:-8}
mkdir -p gpurun_out
run() { local name=$1; shift
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 500)) "$@" > gpurun_out/${name}.json 2> gpurun_out/${name}.err
  echo "$name rc=$?"; cut -c1-1500 gpurun_out/${name}.json; tail -3 gpurun_out/${name}.err | cut -c1-300; }
run j${N}_default bench.py --gpus $N --steps 20 --warmup 3 --sustained 0
ELEPHAS_B200_NVLS=1 run j${N}_default_nvls bench.py --gpus $N --steps 20 --warmup 3 --no-e2e --sustained 0
run j${N}_collective scripts/bench_collective.py
run j${N}_h2d scripts/bench_h2d.py
run j${N}_mnist_async bench.py --gpus $N --config mnist_mlp --mode asynchronous --steps 2000 --warmup 20 --no-e2e
run j${N}_mnist_hogwild bench.py --gpus $N --config mnist_mlp --mode hogwild --steps 2000 --warmup 20 --no-e2e
timeout 300 python -m pytest tests/parallel/test_gpu_spmd.py -m gpu -x -q > gpurun_out/j${N}_spmd.log 2>&1; echo "spmd rc=$?"; tail -30 gpurun_out/j${N}_spmd.log
run j${N}_resnet bench.py --gpus $N --config resnet50
