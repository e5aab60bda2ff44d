#!/bin/bash
mkdir -p gpurun_out
export SPARK_LOCAL_IP=127.0.0.1
bash scripts/gpu_gemm_check.sh > gpurun_out/gemm_sweep_tail.log 2>&1
tail -12 gpurun_out/gemm_sweep_tail.log
bash scripts/gpu_round.sh
