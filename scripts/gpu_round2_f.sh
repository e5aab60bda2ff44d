#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/f_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/f_pytest.log
RESNET_GRAPHS=1 python scripts/resnet_step.py 12 2>&1 | tail -3
RESNET_GRAPHS=1 ELEPHAS_B200_NATIVE_CONV=0 python scripts/resnet_step.py 12 2>&1 | tail -2
python scripts/resnet_step.py 3 > gpurun_out/f_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/f_launches.csv python scripts/resnet_step.py 2 > gpurun_out/f_ncu.log 2>&1
python scripts/summarize_launches.py gpurun_out/f_launches.csv | tee gpurun_out/f_launch_summary.txt | head -30
for mode in asynchronous hogwild; do
  timeout 600 python bench.py --config mnist_mlp --mode $mode --steps 2000 --warmup 20 --no-e2e > gpurun_out/f_mnist_$mode.json 2> gpurun_out/f_mnist_$mode.err; echo "ps $mode rc=$?"; cut -c1-330 gpurun_out/f_mnist_$mode.json
done
