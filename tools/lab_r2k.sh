#!/bin/bash
# ncu --set full of the two product tiles changed in round 2 (inside a fused mul_decompose call), exported on the box.
cd "${GRAFT_REPO_ROOT:-.}"
O=gpurun_out/final5
mkdir -p $O
timeout 150 ncu --set full --clock-control none --import-source on -k regex:matmul_kernel -s 7 -c 1 -f -o /tmp/c3b_mm python tools/md_once.py c3b > $O/full_c3b.log 2>&1; echo "c3b rc=$?"
timeout 150 ncu --set full --clock-control none --import-source on -k regex:matmul_f64_kernel -s 1 -c 1 -f -o /tmp/c4_mm python tools/md_once.py c4 > $O/full_c4.log 2>&1; echo "c4 rc=$?"
ncu -i /tmp/c3b_mm.ncu-rep --page raw --csv > $O/c3b_mm_raw.csv 2>/dev/null
ncu -i /tmp/c4_mm.ncu-rep --page raw --csv > $O/c4_mm_raw.csv 2>/dev/null
python tools/ncu_summary.py $O/c3b_mm_raw.csv; python tools/ncu_summary.py $O/c4_mm_raw.csv
