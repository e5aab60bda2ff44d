#!/bin/bash
# Quick one-GPU check: native-op tests, kernel tour, short bench.
mkdir -p gpurun_out
export SPARK_LOCAL_IP=127.0.0.1
L=gpurun_out/quick.log
: > $L
timeout 400 python -m pytest ${PYTEST_TARGET:-tests/ops} -m gpu -q -x --timeout 90 --timeout-method=thread > gpurun_out/pytest_quick.log 2>&1
echo "pytest rc=$?" >> $L
tail -25 gpurun_out/pytest_quick.log >> $L
timeout 300 python scripts/kernel_tour.py > gpurun_out/kernel_tour.log 2>&1
echo "tour rc=$?" >> $L
cat gpurun_out/kernel_tour.log >> $L
timeout 300 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_quick.json 2> gpurun_out/bench.err
echo "bench rc=$?" >> $L
grep '^{' gpurun_out/bench_quick.json >> $L
tail -3 gpurun_out/bench.err >> $L
cat $L
