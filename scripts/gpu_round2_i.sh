#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/i_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/i_pytest.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/i_bench_default.json 2> gpurun_out/i_bench_default.err; echo "bench rc=$?"; cat gpurun_out/i_bench_default.json; tail -3 gpurun_out/i_bench_default.err
python scripts/bench_h2d.py
timeout 600 python bench.py --config resnet50 > gpurun_out/i_resnet.json 2> gpurun_out/i_resnet.err; cat gpurun_out/i_resnet.json; tail -2 gpurun_out/i_resnet.err
timeout 600 python bench.py --config otto_pipeline > gpurun_out/i_otto.json 2> gpurun_out/i_otto.err; cat gpurun_out/i_otto.json; tail -2 gpurun_out/i_otto.err
