#!/bin/bash
# One-GPU round: gpu tests, smoke, bench A/B (side stream on/off), BASELINE configs 2/3/5, kernel tour + ncu of the
# non-GEMM kernels. Logs land in gpurun_out/.
mkdir -p gpurun_out
export SPARK_LOCAL_IP=127.0.0.1
L=gpurun_out/round3.log
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $L 2>&1
if [ -z "$SKIP_TESTS" ]; then
  echo "=== pytest -m gpu" >> $L
  timeout 480 python -m pytest tests -m gpu -q --timeout 90 --timeout-method=thread > gpurun_out/pytest_gpu.log 2>&1
  echo "pytest rc=$?" >> $L
  tail -15 gpurun_out/pytest_gpu.log >> $L
  echo "=== smoke" >> $L
  timeout 300 python -c "import __graft_entry__ as g; g.smoke()" >> $L 2>&1
  echo "smoke rc=$?" >> $L
fi
for side in 1 0; do
  echo "=== bench side_stream=$side" >> $L
  ELEPHAS_B200_SIDE_STREAM=$side timeout 600 python bench.py --steps ${BENCH_STEPS:-20} --warmup 3 > gpurun_out/bench_side$side.json 2> gpurun_out/bench.err
  echo "bench rc=$?" >> $L
  grep '^{' gpurun_out/bench_side$side.json >> $L
  tail -3 gpurun_out/bench.err >> $L
done
for c in 2 3 5; do
  echo "=== config $c" >> $L
  timeout 600 python benchmarks/configs.py --config $c --gpus 1 > gpurun_out/config$c.out 2> gpurun_out/config$c.err
  echo "rc=$?" >> $L
  grep '^{' gpurun_out/config$c.out | tee gpurun_out/config${c}_1gpu.json >> $L
  tail -3 gpurun_out/config$c.err >> $L
done
echo "=== kernel tour" >> $L
timeout 300 python scripts/kernel_tour.py > gpurun_out/kernel_tour.log 2>&1
rc=$?
echo "tour rc=$rc" >> $L
cat gpurun_out/kernel_tour.log >> $L
if [ $rc -eq 0 ]; then
  TOUR_ONCE=1 timeout 600 ncu --set full --clock-control none --import-source on \
      -k regex:'stage_f32|cast_f32|colsum|act_dropout|delta_kernel|loss_kernel|optimizer_kernel|ps_push|ps_pull|allreduce' \
      --launch-count 40 -f -o gpurun_out/kernel_tour_prof python scripts/kernel_tour.py > gpurun_out/ncu_tour.log 2>&1
  echo "ncu rc=$?" >> $L
  ncu -i gpurun_out/kernel_tour_prof.ncu-rep --page raw --csv \
      --metrics gpu__time_duration.sum,dram__throughput.avg.pct_of_peak_sustained_elapsed,dram__bytes.sum,sm__throughput.avg.pct_of_peak_sustained_elapsed,launch__registers_per_thread,launch__grid_size,launch__block_size,sm__warps_active.avg.pct_of_peak_sustained_active \
      > gpurun_out/kernel_tour_ncu.csv 2>> $L
fi
tail -80 $L
