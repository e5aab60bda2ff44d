#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/g_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/g_pytest.log
RESNET_GRAPHS=1 python scripts/resnet_step.py 12 2>&1 | tail -2
python scripts/resnet_step.py 3 > gpurun_out/g_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/g_launches.csv python scripts/resnet_step.py 2 > gpurun_out/g_ncu.log 2>&1
python scripts/summarize_launches.py gpurun_out/g_launches.csv | tee gpurun_out/g_launch_summary.txt | head -24
