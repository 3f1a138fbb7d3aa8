#!/bin/bash
# Four-row product tile A/B inside the fused mul_decompose, the no-alloc test alone in a fresh process, then the whole GPU suite.
cd "${GRAFT_REPO_ROOT:-.}"
O=gpurun_out/lab7
mkdir -p $O
for o in 0 1; do
  MXX_MATMUL_ROWS4=$o timeout 200 python tools/md_lanes_ab.py > $O/md_rows4_$o.log 2>&1; echo "rows4=$o rc=$?"; grep "C3-b\|C3-a\|28-bit" $O/md_rows4_$o.log
done
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x > $O/pytest_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; tail -4 $O/pytest_noalloc_alone.log
timeout 900 python -m pytest tests -m gpu -q -x --durations=10 > $O/pytest_gpu_all.log 2>&1; echo "gpu suite rc=$?"; tail -16 $O/pytest_gpu_all.log
