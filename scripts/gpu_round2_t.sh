#!/bin/bash
# round 2, GPU call T (N GPUs): multi-rank correctness script, parameter-server modes (counter / gate protocol), ResNet-50, headline
# line and the collective benchmark after the last kernel changes
set -x
N=${1:-4}
mkdir -p gpurun_out
timeout 300 python -m pytest tests/parallel/test_gpu_spmd.py -m gpu -x -q > gpurun_out/t${N}_spmd.log 2>&1; echo "spmd rc=$?"; tail -20 gpurun_out/t${N}_spmd.log
run() { local name=$1; shift
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 500)) "$@" > gpurun_out/${name}.json 2> gpurun_out/${name}.err
  echo "$name rc=$?"; cut -c1-1200 gpurun_out/${name}.json; tail -3 gpurun_out/${name}.err | cut -c1-300; }
run t${N}_default bench.py --gpus $N --steps 20 --warmup 3 --sustained 0
run t${N}_mnist_async bench.py --gpus $N --config mnist_mlp --mode asynchronous --steps 2000 --warmup 20 --no-e2e
run t${N}_mnist_hogwild bench.py --gpus $N --config mnist_mlp --mode hogwild --steps 2000 --warmup 20 --no-e2e
run t${N}_resnet bench.py --gpus $N --config resnet50
run t${N}_collective scripts/bench_collective.py
