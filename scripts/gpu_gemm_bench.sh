#!/bin/bash
T=build/test_gemm
for shape in "4096 4096 4096" "8192 8192 8192" "4096 4096 8192"; do
for cfg in "0 0 256 2" "0 1 256 2" "1 1 256 2"; do
  timeout 60 $T bench $shape $cfg 10
done; done
timeout 60 $T check 512 512 512 0 0 256 2
timeout 60 $T check 1000 777 520 1 1 256 2 0 0 0 2
timeout 60 $T check 4096 4096 1024 0 0 0 0 0 1 0.2
