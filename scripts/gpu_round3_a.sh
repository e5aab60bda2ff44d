#!/bin/bash
# First GPU run of the Keras-surface routes added at the end of round 2 (tests/ops/test_keras_surface_gpu.py).
# /usr/local/graft/bin/gpurun --timeout 900 -- 'bash scripts/gpu_round3_a.sh'
set -u
mkdir -p gpurun_out
ELEPHAS_B200_PENDING_GPU_TESTS=1 timeout 600 python -m pytest tests/ops/test_keras_surface_gpu.py -m gpu -q -x \
    > gpurun_out/r3a_keras_surface.log 2>&1
echo "keras surface tests exit $?" | tee -a gpurun_out/r3a_keras_surface.log
tail -15 gpurun_out/r3a_keras_surface.log
python -m elephas_b200 > gpurun_out/r3a_info.txt 2>&1
cat gpurun_out/r3a_info.txt
