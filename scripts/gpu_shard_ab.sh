#!/bin/bash
# A/B of the sharded-master fused all-reduce+update against the replicated one on N GPUs (SPMD checks first).
N=${1:-2}
mkdir -p gpurun_out
export SPARK_LOCAL_IP=127.0.0.1
L=gpurun_out/shard_ab_$N.log
nvidia-smi --query-gpu=index,name --format=csv > $L 2>&1
echo "=== spmd checks" >> $L
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 \
    tests/parallel/gpu_spmd_script.py 7400 >> $L 2>&1
echo "spmd rc=$?" >> $L
for sharded in 1 0; do
  for blocks in ${COMM_BLOCKS_LIST:-74}; do
    echo "=== bench ours N=$N sharded=$sharded blocks=$blocks" >> $L
    ELEPHAS_B200_SHARDED_MASTER=$sharded ELEPHAS_B200_COMM_BLOCKS=$blocks timeout 400 python -m torch.distributed.run --nnodes=1 \
        --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29612 \
        bench.py --gpus $N --steps ${BENCH_STEPS:-20} --warmup 3 > gpurun_out/ab.json 2> gpurun_out/ab.err
    echo "bench rc=$?" >> $L
    grep '^{' gpurun_out/ab.json >> $L
    tail -3 gpurun_out/ab.err >> $L
  done
done
tail -40 $L
