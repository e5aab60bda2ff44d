#!/bin/bash
# round 2, GPU call D (1 GPU): K9 numerics, then ResNet-50 native vs cuDNN
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/ops/test_conv_gpu.py -m gpu -q -x > gpurun_out/d_conv_tests.log 2>&1; rc=$?; echo "conv tests rc=$rc"; tail -40 gpurun_out/d_conv_tests.log
if [ $rc -eq 0 ]; then
  for flag in 1 0; do
    ELEPHAS_B200_NATIVE_CONV=$flag timeout 900 python bench.py --config resnet50 > gpurun_out/d_resnet_native$flag.json 2> gpurun_out/d_resnet_native$flag.err; echo "resnet native=$flag rc=$?"
    cat gpurun_out/d_resnet_native$flag.json; tail -5 gpurun_out/d_resnet_native$flag.err
  done
fi
