#!/bin/bash
# Launch list (per-kernel device time) of a short bench run; graphs off so every kernel is a plain launch.
mkdir -p gpurun_out
export ELEPHAS_B200_GRAPHS=0
CMD="python bench.py --steps 2 --warmup 1 --no-e2e"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/plain.log; wc -l gpurun_out/launches.csv
