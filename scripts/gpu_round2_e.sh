#!/bin/bash
set -x
mkdir -p gpurun_out
python scripts/resnet_step.py 3 > gpurun_out/e_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/e_launches.csv python scripts/resnet_step.py 2 > gpurun_out/e_ncu.log 2>&1
cat gpurun_out/e_plain.log | tail -5
python scripts/summarize_launches.py gpurun_out/e_launches.csv | tee gpurun_out/e_launch_summary.txt | head -50
RESNET_GRAPHS=1 python scripts/resnet_step.py 12 2>&1 | tail -4
