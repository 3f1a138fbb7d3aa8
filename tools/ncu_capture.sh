#!/bin/bash
# Runs on the GPU box (under gpurun): one `ncu --set full` capture per hot kernel, exported to CSV on the box so
# that only small text files come back in gpurun_out/ (the .ncu-rep files stay in /tmp unless small).
# usage: tools/ncu_capture.sh <tag> [what ...]   (what = ntt32 ntt64 ntt16 md mul)
set -u
export PROF_REPS=1
tag=$1; shift
mkdir -p gpurun_out
for what in "$@"; do
  python tools/prof_kernels.py $what > gpurun_out/plain_${what}.log 2>&1 || { echo "plain run of $what failed"; continue; }
  rep=/tmp/${tag}_${what}
  ncu --set full --clock-control none --import-source on -k regex:"fast_|matmul_kernel|matmul_f64|global_pass|decompose|mac" -s ${SKIP:-0} -c ${COUNT:-6} \
      -o $rep -f python tools/prof_kernels.py $what > gpurun_out/ncu_${what}.log 2>&1
  ncu -i $rep.ncu-rep --page raw --csv > gpurun_out/${tag}_${what}_raw.csv 2>/dev/null
  ncu -i $rep.ncu-rep --page source --csv > gpurun_out/${tag}_${what}_source.csv 2>/dev/null
  sz=$(stat -c %s $rep.ncu-rep 2>/dev/null || echo 0)
  if [ "$sz" -lt 12000000 ]; then cp $rep.ncu-rep gpurun_out/; fi
done
