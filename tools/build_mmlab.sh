#!/bin/bash
mkdir -p lab
for spec in "$@"; do
  tag=${spec%%:*}; flags=${spec#*:}
  ( /usr/local/cuda/bin/nvcc -std=c++17 -O3 -lineinfo -gencode arch=compute_100a,code=sm_100a -Imxx_b200/csrc/include -cudart static $flags -o lab/mm_$tag tools/matmul_lab.cu 2>&1 | grep -E "error" | head -3 ) &
done
wait
