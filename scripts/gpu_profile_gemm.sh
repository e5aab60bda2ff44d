#!/bin/bash
# ncu --set full capture of the GEMM kernel (one GPU, tiny command), per B200_PROFILING.md.
mkdir -p gpurun_out
CMD="build/test_gemm bench ${GEMM_SHAPE:-4096 4096 4096} 0 0 256 2 3"
$CMD > gpurun_out/plain_gemm.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:gemm_bf16 -s 3 -c 1 -f -o gpurun_out/gemm_prof $CMD > gpurun_out/ncu_gemm.log 2>&1
echo "rc=$?"; cat gpurun_out/plain_gemm.log; tail -3 gpurun_out/ncu_gemm.log; ls -la gpurun_out/*.ncu-rep
