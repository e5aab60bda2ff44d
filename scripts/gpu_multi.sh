#!/bin/bash
# Multi-GPU round: SPMD checks + scaling bench (ours and DDP) on N GPUs of one box.
N=${1:-2}
mkdir -p gpurun_out
export SPARK_LOCAL_IP=127.0.0.1
nvidia-smi --query-gpu=index,name --format=csv > gpurun_out/multi_$N.log 2>&1
echo "=== spmd checks" >> gpurun_out/multi_$N.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 \
    tests/parallel/gpu_spmd_script.py 7400 >> gpurun_out/multi_$N.log 2>&1
echo "spmd rc=$?" >> gpurun_out/multi_$N.log
for impl in ours ddp; do
  echo "=== bench $impl N=$N" >> gpurun_out/multi_$N.log
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29612 \
      bench.py --gpus $N --steps ${BENCH_STEPS:-20} --warmup 3 --impl $impl > gpurun_out/bench_${impl}_$N.json 2> gpurun_out/bench_${impl}_$N.err
  echo "bench rc=$?" >> gpurun_out/multi_$N.log
  grep '^{' gpurun_out/bench_${impl}_$N.json >> gpurun_out/multi_$N.log
  tail -5 gpurun_out/bench_${impl}_$N.err >> gpurun_out/multi_$N.log
done
echo "=== bench ddp N=1" >> gpurun_out/multi_$N.log
CUDA_VISIBLE_DEVICES=0 timeout 300 python bench.py --gpus 1 --steps 20 --warmup 3 --impl ddp 2>/dev/null | grep '^{' >> gpurun_out/multi_$N.log
tail -60 gpurun_out/multi_$N.log
