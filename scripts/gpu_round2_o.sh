#!/bin/bash
# round 2, GPU call O (1 GPU): GPU test suite, smoke(), and the one-GPU column of the matrix (ours + torch arms)
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/o_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/o_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
run() { local name=$1; shift
  timeout 300 python "$@" > gpurun_out/${name}.json 2> gpurun_out/${name}.err
  echo "$name rc=$?"; cut -c1-1200 gpurun_out/${name}.json; tail -2 gpurun_out/${name}.err | cut -c1-300; }
run o1_default bench.py --steps 20 --warmup 3
run o1_ddp_wide bench.py --impl ddp --steps 20 --warmup 3
run o1_mnist_async bench.py --config mnist_mlp --mode asynchronous --steps 2000 --warmup 20
run o1_mnist_hogwild bench.py --config mnist_mlp --mode hogwild --steps 2000 --warmup 20
run o1_ddp_mnist bench.py --impl ddp --config mnist_mlp --steps 2000 --warmup 20
run o1_ddp_resnet bench.py --impl ddp --config resnet50
run o1_resnet bench.py --config resnet50
ELEPHAS_B200_NATIVE_CONV=0 run o1_resnet_cudnn bench.py --config resnet50
run o1_otto bench.py --config otto_pipeline
run o1_reference bench.py --impl reference
