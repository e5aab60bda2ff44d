#!/bin/bash
# End-to-end (SparkModel.fit, streaming inputs) phase breakdown on N GPUs: tracing on, two step counts.
N=${1:-2}
mkdir -p gpurun_out
export SPARK_LOCAL_IP=127.0.0.1
L=gpurun_out/e2e_trace_$N.log
: > $L
for K in ${STEPS_LIST:-20 60}; do
  echo "=== N=$N steps=$K" >> $L
  ELEPHAS_B200_TRACE=1 timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
      --master-port 29612 bench.py --gpus $N --steps $K --warmup 3 > gpurun_out/tr.json 2> gpurun_out/tr.err
  echo "rc=$?" >> $L
  grep '^{' gpurun_out/tr.json >> $L
  grep -A30 "^host " gpurun_out/tr.err >> $L
  grep "staging choices" gpurun_out/tr.err >> $L
  grep -i "error\|Traceback" gpurun_out/tr.err | head -3 >> $L
done
cat $L
