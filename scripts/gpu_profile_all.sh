#!/bin/bash
# Evidence run: ncu --set full of the top kernel (forward Dense GEMM, 4096^3, CTA pair) inside the bench
# process, plus the launch list of one step. One GPU; graphs off so kernels are plain launches.
mkdir -p gpurun_out
export ELEPHAS_B200_GRAPHS=0
CMD="python bench.py --steps 2 --warmup 1 --no-e2e"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:gemm_bf16_kernel -s 30 -c 3 -f -o gpurun_out/step_gemm_prof $CMD > gpurun_out/ncu_full.log 2>&1
echo "rc=$?"; tail -2 gpurun_out/ncu_full.log; ls -la gpurun_out/step_gemm_prof.ncu-rep
