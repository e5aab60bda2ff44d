#!/bin/bash
# round 2, GPU call R (N GPUs): parameter-server modes after the protocol change, ResNet-50 (ours and torchvision DDP), the collective
# benchmark with the NCCL arm, the headline line
set -x
N=${1:-8}
mkdir -p gpurun_out
run() { local name=$1; shift
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 500)) "$@" > gpurun_out/${name}.json 2> gpurun_out/${name}.err
  echo "$name rc=$?"; cut -c1-1200 gpurun_out/${name}.json; tail -3 gpurun_out/${name}.err | cut -c1-300; }
run r${N}_mnist_async bench.py --gpus $N --config mnist_mlp --mode asynchronous --steps 2000 --warmup 20 --no-e2e
run r${N}_mnist_hogwild bench.py --gpus $N --config mnist_mlp --mode hogwild --steps 2000 --warmup 20 --no-e2e
run r${N}_resnet bench.py --gpus $N --config resnet50
run r${N}_ddp_resnet bench.py --gpus $N --impl ddp --config resnet50
run r${N}_collective scripts/bench_collective.py
run r${N}_default bench.py --gpus $N --steps 20 --warmup 3 --sustained 0
