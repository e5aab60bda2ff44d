#!/bin/bash
# Numerics sweep + speed sweep of the tcgen05 GEMM on a B200 (run through gpurun).
# Every case runs in its own process under `timeout` so a hang cannot take the sweep down.
mkdir -p gpurun_out
LOG=gpurun_out/gemm_check.log
: > $LOG
T=build/test_gemm
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv >> $LOG 2>&1
run() { timeout 40 $T "$@" >> $LOG 2>&1; rc=$?; if [ $rc -ne 0 ]; then echo "RC=$rc for: $*" >> $LOG; fi; }
# --- 1-CTA, K-major
run check 128 128 64 0 0 128 1
run check 256 256 256 0 0 128 1
run check 200 136 200 0 0 64 1
run check 64 10 784 0 0 16 1 0 2 0
run check 100 40 100 0 0 32 1 0 1 0
run check 512 512 512 0 0 256 1
run check 1024 768 2048 0 0 256 1
# --- MN-major operands
run check 256 256 256 1 0 128 1
run check 256 256 256 0 1 128 1
run check 256 256 256 1 1 128 1
run check 200 136 200 1 1 64 1
run check 200 136 200 0 1 64 1
run check 200 136 200 1 0 64 1
run check 10 128 64 1 1 64 1
run check 512 512 512 1 1 256 1
# --- CTA pairs
run check 256 128 64 0 0 128 2
run check 512 512 512 0 0 256 2
run check 512 256 512 0 0 128 2
run check 512 512 512 1 1 256 2
run check 512 512 512 0 1 256 2
run check 512 512 512 1 0 128 2
run check 300 200 200 0 0 128 2
run check 300 200 200 1 1 128 2
# --- epilogues
run check 256 256 256 0 0 128 1 0 1 0.2
run check 256 256 256 0 0 128 1 1 1 0.2
run check 256 256 256 0 1 128 1 1 2 0.0
run check 256 256 256 0 0 128 1 1 3 0.5
run check 256 256 256 1 1 128 1 0 0 0 1
run check 256 256 256 1 1 128 1 0 0 0 2
run check 200 136 200 1 1 64 1 0 0 0 2
run check 512 512 512 1 1 256 2 0 0 0 2
# --- persistent multi-tile, auto config
run check 4096 4096 1024 0 0 0 0
run check 4096 4096 1024 0 1 0 0 1 1 0
run check 4096 4096 1024 1 1 0 0 0 0 0 2
run check 2048 3000 520 0 0 0 0 0 1 0.2
# --- split tail wave (tiles % CTA-groups small): 256 tiles on 74 pairs, 160 tiles on 148 CTAs
run check 4096 4096 512 0 0 256 2
run check 4096 4096 512 0 1 256 2 1 1 0.2
run check 4096 4096 512 1 1 256 2 0 0 0 2
run check 4096 4000 520 0 0 256 2 0 1 0.2
run check 1280 4096 256 0 0 256 1 0 2 0
run check 1280 4096 256 1 1 256 1 0 0 0 1
run check 1280 4090 256 0 1 256 1 1 3 0
# --- speed (L2 flushed between iterations, operands > L2)
for cfg in "256 1" "256 2" "128 1" "128 2"; do
  run bench 8192 8192 8192 0 0 $cfg 10
done
run bench 8192 8192 8192 0 1 256 2 10
run bench 8192 8192 8192 1 1 256 2 10
run bench 4096 4096 4096 0 0 256 2 10
run bench 4096 4096 4096 0 0 256 1 10
run bench 4096 4096 4096 0 0 0 0 10
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv >> $LOG 2>&1
echo DONE >> $LOG
grep -c PASS $LOG; grep -c FAIL $LOG; grep RC= $LOG | head -20; grep BENCH $LOG
