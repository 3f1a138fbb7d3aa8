#!/bin/bash
# builds lab/mmr_<tag> for each "tag:flags" argument (development only; lab/ is git-ignored but travels to the GPU box)
mkdir -p lab
for spec in "$@"; do
  tag=${spec%%:*}; flags=${spec#*:}
  ( /usr/local/cuda/bin/nvcc -std=c++17 -O3 -lineinfo -gencode arch=compute_100a,code=sm_100a -Imxx_b200/csrc/include -cudart static $flags -o lab/mmr_$tag tools/matmul_rows_lab.cu 2>&1 | grep -E "error" | head -3 ) &
done
wait
