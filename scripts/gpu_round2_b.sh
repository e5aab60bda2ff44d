#!/bin/bash
# round 2, GPU call B (2 GPUs): multi-rank correctness script, scaling of the PS configs and of the default config, torch arms
set -x
N=${1:-2}
mkdir -p gpurun_out
run() { # name, args...
  local name=$1; shift
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 500)) "$@" > gpurun_out/${name}.json 2> gpurun_out/${name}.err
  echo "$name rc=$?"; cat gpurun_out/${name}.json; tail -4 gpurun_out/${name}.err
}
timeout 900 python -m pytest tests/parallel/test_gpu_spmd.py -m gpu -x -q > gpurun_out/b${N}_spmd.log 2>&1; echo "spmd rc=$?"; tail -30 gpurun_out/b${N}_spmd.log
run b${N}_default bench.py --gpus $N --steps 20 --warmup 3
run b${N}_mnist_async bench.py --gpus $N --config mnist_mlp --mode asynchronous --steps 2000 --warmup 20
run b${N}_mnist_hogwild bench.py --gpus $N --config mnist_mlp --mode hogwild --steps 2000 --warmup 20
run b${N}_ddp_wide bench.py --gpus $N --impl ddp --steps 20 --warmup 3
run b${N}_ddp_mnist bench.py --gpus $N --impl ddp --config mnist_mlp --steps 2000 --warmup 20
run b${N}_localsgd_wide bench.py --gpus $N --impl localsgd --steps 20 --warmup 3
