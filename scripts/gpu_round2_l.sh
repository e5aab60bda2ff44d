#!/bin/bash
# round 2, GPU call L (1 GPU): full GPU test suite, configs 3 and 5 through bench.py, and full ncu captures of the convolution /
# BatchNorm kernels inside an eager ResNet-50 step (the plain run first, ncu only if it exits 0)
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/l_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/l_pytest.log
timeout 300 python bench.py --config resnet50 > gpurun_out/l_resnet.json 2> gpurun_out/l_resnet.err; cat gpurun_out/l_resnet.json; tail -2 gpurun_out/l_resnet.err
timeout 300 python bench.py --config otto_pipeline > gpurun_out/l_otto.json 2> gpurun_out/l_otto.err; cat gpurun_out/l_otto.json; tail -2 gpurun_out/l_otto.err
python scripts/resnet_step.py 3 > gpurun_out/l_plain.log 2>&1 && {
for k in conv_igemm_kernel conv_wgrad_kernel; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -s 40 -c 3 -f -o gpurun_out/l_$k python scripts/resnet_step.py 3 > gpurun_out/l_ncu_$k.log 2>&1; echo "ncu $k rc=$?"
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'stats_kernel|bwd_reduce_kernel|bwd_apply_kernel|apply_kernel' -s 60 -c 4 -f -o gpurun_out/l_bn python scripts/resnet_step.py 3 > gpurun_out/l_ncu_bn.log 2>&1; echo "ncu bn rc=$?"
}
tail -3 gpurun_out/l_plain.log; ls -la gpurun_out/*.ncu-rep
