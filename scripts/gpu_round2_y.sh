#!/bin/bash
# round 2, GPU call Y (1 GPU): ncu --set full of the parameter-server kernels inside the asynchronous per-batch loop (plain run first)
set -x
mkdir -p gpurun_out
CMD="python bench.py --config mnist_mlp --mode asynchronous --steps 60 --warmup 5 --no-e2e"
ELEPHAS_B200_GRAPHS=0 $CMD > gpurun_out/y_plain.log 2>&1 && \
ELEPHAS_B200_GRAPHS=0 timeout 300 ncu --set full --clock-control none --import-source on -k regex:'ps_pull_kernel|ps_optimizer_push_kernel' -s 20 -c 4 -f -o gpurun_out/y_ps $CMD > gpurun_out/y_ncu.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/y_plain.log | cut -c1-300; tail -5 gpurun_out/y_ncu.log | cut -c1-200; ls -la gpurun_out/y_ps.ncu-rep
