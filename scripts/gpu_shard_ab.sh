#!/bin/bash
# A/B of the sharded-master fused all-reduce+update against the replicated one on N GPUs (SPMD checks first).
# VARIANTS: space-separated "sharded:blocks[:skip]" triples.
N=${1:-2}
mkdir -p gpurun_out
export SPARK_LOCAL_IP=127.0.0.1
L=gpurun_out/shard_ab_$N.log
nvidia-smi --query-gpu=index,name --format=csv > $L 2>&1
echo "=== spmd checks" >> $L
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 \
    tests/parallel/gpu_spmd_script.py 7400 > gpurun_out/spmd_$N.log 2>&1
echo "spmd rc=$?" >> $L
grep "ok\|OK\|Error\|error" gpurun_out/spmd_$N.log | tail -25 >> $L
for v in ${VARIANTS:-1:74 0:74}; do
  IFS=: read sharded blocks skip <<< "$v"
  echo "=== bench ours N=$N sharded=$sharded blocks=$blocks skip_comm=${skip:-0}" >> $L
  ELEPHAS_B200_DEBUG_SKIP_COMM=${skip:-0} ELEPHAS_B200_SHARDED_MASTER=$sharded ELEPHAS_B200_COMM_BLOCKS=$blocks timeout 400 \
      python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29612 \
      bench.py --gpus $N --steps ${BENCH_STEPS:-20} --warmup 3 > gpurun_out/ab.json 2> gpurun_out/ab.err
  echo "bench rc=$?" >> $L
  python - >> $L <<'PY'
import json
for line in open('gpurun_out/ab.json'):
    if line.startswith('{'):
        d = json.loads(line)
        print("ms_per_step %.4f  value %.0f  e2e_ms %.4f  clocks %s" % (d['ms_per_step'], d['value'], d['e2e']['ms_per_step'], d['clocks']))
PY
  grep -i "error\|Traceback" gpurun_out/ab.err | head -3 >> $L
done
cat $L
