#!/bin/bash
# One-GPU evidence round: gpu tests, smoke, bench, BASELINE configs 2/3/5, kernel tour (+ncu), in-step GEMM ncu,
# compute-sanitizer memcheck of the small kernels. Everything lands in gpurun_out/.
mkdir -p gpurun_out
export SPARK_LOCAL_IP=127.0.0.1
L=gpurun_out/final.log
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $L 2>&1
echo "=== pytest -m gpu" >> $L
timeout 600 python -m pytest tests -m gpu -q --timeout 120 --timeout-method=thread > gpurun_out/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> $L
tail -8 gpurun_out/pytest_gpu.log >> $L
echo "=== smoke" >> $L
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" >> $L 2>&1
echo "smoke rc=$?" >> $L
echo "=== bench" >> $L
timeout 600 python bench.py --steps ${BENCH_STEPS:-30} --warmup 3 > gpurun_out/bench_final.json 2> gpurun_out/bench.err
echo "bench rc=$?" >> $L
grep '^{' gpurun_out/bench_final.json >> $L
tail -3 gpurun_out/bench.err >> $L
for c in 2 3 5; do
  echo "=== config $c" >> $L
  timeout 600 python benchmarks/configs.py --config $c --gpus 1 > gpurun_out/config$c.out 2> gpurun_out/config$c.err
  echo "rc=$?" >> $L
  grep '^{' gpurun_out/config$c.out | tee gpurun_out/config${c}_1gpu.json >> $L
  tail -2 gpurun_out/config$c.err >> $L
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
  echo "ncu tour rc=$?" >> $L
fi
echo "=== in-step GEMM ncu" >> $L
export ELEPHAS_B200_GRAPHS=0
CMD="python bench.py --steps 2 --warmup 1 --no-e2e"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_bf16_kernel -s 30 -c 4 -f -o gpurun_out/step_gemm_prof $CMD > gpurun_out/ncu_gemm.log 2>&1
echo "ncu gemm rc=$?" >> $L
unset ELEPHAS_B200_GRAPHS
echo "=== compute-sanitizer memcheck (small kernels)" >> $L
timeout 420 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/ops/test_native_ops.py -m gpu -q -x \
    -k "loss or optimizers or stage_cast or vector_and_scalar or parameter_server or allreduce_update_single or emulated or delta_mode" \
    > gpurun_out/sanitizer_memcheck.log 2>&1
echo "memcheck rc=$?" >> $L
tail -6 gpurun_out/sanitizer_memcheck.log >> $L
tail -70 $L
