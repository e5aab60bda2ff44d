#!/bin/bash
# round 2, GPU call M (N GPUs): the whole matrix at N - ours (headline with e2e, parameter-server modes, ResNet-50), the torch DDP
# and local-SGD arms, collective roofline, host-link ceiling, multi-rank correctness script
set -x
N=${1:-4}
mkdir -p gpurun_out
run() { local name=$1; shift
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 500)) "$@" > gpurun_out/${name}.json 2> gpurun_out/${name}.err
  echo "$name rc=$?"; cut -c1-1200 gpurun_out/${name}.json; tail -3 gpurun_out/${name}.err | cut -c1-300; }
run m${N}_default bench.py --gpus $N --steps 20 --warmup 3
run m${N}_ddp_wide bench.py --gpus $N --impl ddp --steps 20 --warmup 3
run m${N}_localsgd_wide bench.py --gpus $N --impl localsgd --steps 20 --warmup 3
run m${N}_mnist_async bench.py --gpus $N --config mnist_mlp --mode asynchronous --steps 2000 --warmup 20 --no-e2e
run m${N}_mnist_hogwild bench.py --gpus $N --config mnist_mlp --mode hogwild --steps 2000 --warmup 20 --no-e2e
run m${N}_ddp_mnist bench.py --gpus $N --impl ddp --config mnist_mlp --steps 2000 --warmup 20
run m${N}_collective scripts/bench_collective.py
run m${N}_h2d scripts/bench_h2d.py
run m${N}_resnet bench.py --gpus $N --config resnet50
run m${N}_ddp_resnet bench.py --gpus $N --impl ddp --config resnet50
timeout 300 python -m pytest tests/parallel/test_gpu_spmd.py -m gpu -x -q > gpurun_out/m${N}_spmd.log 2>&1; echo "spmd rc=$?"; tail -30 gpurun_out/m${N}_spmd.log
