#!/bin/bash
# round 2, GPU call A (1 GPU): test-suite, default bench, parameter-server configs, torch baselines
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/a_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/a_pytest.log
tail -15 gpurun_out/a_pytest.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/a_bench_default.json 2> gpurun_out/a_bench_default.err; echo "bench rc=$?"
cat gpurun_out/a_bench_default.json; tail -5 gpurun_out/a_bench_default.err
for mode in asynchronous hogwild; do
  timeout 600 python bench.py --config mnist_mlp --mode $mode --steps 2000 --warmup 20 > gpurun_out/a_bench_mnist_$mode.json 2> gpurun_out/a_bench_mnist_$mode.err; echo "ps $mode rc=$?"
  cat gpurun_out/a_bench_mnist_$mode.json; tail -5 gpurun_out/a_bench_mnist_$mode.err
done
timeout 600 python bench.py --impl ddp --steps 20 --warmup 3 > gpurun_out/a_ddp_wide.json 2> gpurun_out/a_ddp_wide.err; cat gpurun_out/a_ddp_wide.json; tail -3 gpurun_out/a_ddp_wide.err
timeout 600 python bench.py --impl ddp --config mnist_mlp --steps 2000 --warmup 20 > gpurun_out/a_ddp_mnist.json 2> gpurun_out/a_ddp_mnist.err; cat gpurun_out/a_ddp_mnist.json; tail -3 gpurun_out/a_ddp_mnist.err
