#!/bin/bash
# round 2, GPU call Q (N GPUs): whole GPU suite (incl. the multi-rank script) and the parameter-server modes after the protocol change
set -x
N=${1:-2}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/q${N}_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/q${N}_pytest.log
run() { local name=$1; shift
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 500)) "$@" > gpurun_out/${name}.json 2> gpurun_out/${name}.err
  echo "$name rc=$?"; cut -c1-900 gpurun_out/${name}.json; tail -3 gpurun_out/${name}.err | cut -c1-300; }
run q${N}_mnist_async bench.py --gpus $N --config mnist_mlp --mode asynchronous --steps 2000 --warmup 20 --no-e2e
run q${N}_mnist_hogwild bench.py --gpus $N --config mnist_mlp --mode hogwild --steps 2000 --warmup 20 --no-e2e
timeout 200 python bench.py --config mnist_mlp --mode asynchronous --steps 2000 --warmup 20 --no-e2e > gpurun_out/q1_mnist_async.json 2> gpurun_out/q1_mnist_async.err; cut -c1-600 gpurun_out/q1_mnist_async.json
