#!/bin/bash
# One GPU round trip: gpu tests, smoke, short bench. Logs land in gpurun_out/.
mkdir -p gpurun_out
export SPARK_LOCAL_IP=127.0.0.1
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/round.log 2>&1
echo "=== pytest -m gpu" >> gpurun_out/round.log
timeout 420 python -m pytest tests -m gpu -q --timeout 90 --timeout-method=thread ${PYTEST_ARGS} > gpurun_out/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> gpurun_out/round.log
tail -40 gpurun_out/pytest_gpu.log >> gpurun_out/round.log
echo "=== smoke" >> gpurun_out/round.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" >> gpurun_out/round.log 2>&1
echo "smoke rc=$?" >> gpurun_out/round.log
echo "=== bench" >> gpurun_out/round.log
timeout 600 python bench.py --steps ${BENCH_STEPS:-10} --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err
echo "bench rc=$?" >> gpurun_out/round.log
cat gpurun_out/bench.json >> gpurun_out/round.log
tail -20 gpurun_out/bench.err >> gpurun_out/round.log
tail -60 gpurun_out/round.log
