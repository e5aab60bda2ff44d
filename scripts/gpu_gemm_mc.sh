#!/bin/bash
T=build/test_gemm
export EB_MC=2
for args in "512 256 64 0 0 256 2" "1024 512 512 0 0 256 2" "1024 512 512 0 1 256 2" "1024 512 512 1 1 256 2" "1024 512 512 1 0 256 2" \
            "4096 4096 512 0 0 256 2" "4096 4096 512 0 1 256 2 1 1 0.2" "4096 4096 512 1 1 256 2 0 0 0 2" "1000 777 520 0 0 256 2 0 1 0.2" "300 200 200 1 1 256 2"; do
  timeout 60 $T check $args; echo "rc=$?"
done
for shape in "4096 4096 4096" "8192 8192 8192"; do
  for cfg in "0 0 256 2" "0 1 256 2" "1 1 256 2"; do timeout 60 $T bench $shape $cfg 10; done
done
export EB_MC=1
timeout 60 $T bench 4096 4096 4096 0 0 256 2 10
timeout 60 $T bench 8192 8192 8192 0 0 256 2 10
