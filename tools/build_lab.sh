#!/bin/bash
# builds lab/ntt_lab_<tag> for each "tag:flags" argument (development only; lab/ is git-ignored)
mkdir -p lab
for spec in "$@"; do
  tag=${spec%%:*}; flags=${spec#*:}
  /usr/local/cuda/bin/nvcc -std=c++17 -O3 -lineinfo -gencode arch=compute_100a,code=sm_100a -Imxx_b200/csrc/include -cudart static $flags -o lab/ntt_lab_$tag tools/ntt_lab.cu 2>&1 | grep -v deprecated &
done
wait
ls -la lab/
