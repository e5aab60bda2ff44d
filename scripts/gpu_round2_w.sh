#!/bin/bash
# round 2, GPU call W (1 GPU): final regression - GPU suite, smoke(), headline line, loss parity over 200 steps (ours vs torch DDP arm)
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/w_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/w_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 300 python bench.py > gpurun_out/w_default.json 2> gpurun_out/w_default.err; echo "bench rc=$?"; cut -c1-1800 gpurun_out/w_default.json
timeout 300 python bench.py --steps 200 --warmup 3 --sustained 0 --no-e2e > gpurun_out/w_ours_200.json 2> gpurun_out/w_ours_200.err; cut -c1-700 gpurun_out/w_ours_200.json
timeout 300 python bench.py --impl ddp --steps 200 --warmup 3 > gpurun_out/w_ddp_200.json 2> gpurun_out/w_ddp_200.err; cut -c1-700 gpurun_out/w_ddp_200.json
