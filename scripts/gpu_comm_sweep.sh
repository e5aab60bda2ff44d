#!/bin/bash
# Sweep of the number of CTAs the overlapped all-reduce kernels may use (N GPUs).
N=${1:-2}
mkdir -p gpurun_out
export SPARK_LOCAL_IP=127.0.0.1
: > gpurun_out/comm_sweep_$N.log
for blocks in 32 74 148; do
  echo "=== COMM_BLOCKS=$blocks" >> gpurun_out/comm_sweep_$N.log
  ELEPHAS_B200_COMM_BLOCKS=$blocks timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29613 \
      bench.py --gpus $N --steps 20 --warmup 3 --no-e2e 2>/dev/null | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['value'], d['gpu_launches'])" >> gpurun_out/comm_sweep_$N.log
done
echo "=== ours N=1" >> gpurun_out/comm_sweep_$N.log
CUDA_VISIBLE_DEVICES=0 timeout 300 python bench.py --gpus 1 --steps 20 --warmup 3 --no-e2e 2>/dev/null | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['value'], d['gpu_launches'])" >> gpurun_out/comm_sweep_$N.log
cat gpurun_out/comm_sweep_$N.log
