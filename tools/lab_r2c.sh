#!/bin/bash
# Round-2 lab call (one GPU): sampler kernels (chained word-per-trip loop, K tasks per lane) and the launch order of the
# row-tiled products.  Writes gpurun_out/lab3/.  gpurun --timeout 720 -- 'bash tools/lab_r2c.sh'
cd "${GRAFT_REPO_ROOT:-.}"
O=gpurun_out/lab3
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > $O/gpu.txt
t0=$(date +%s)
for k in 1 2 4 8; do
  MXX_SAMPLER_BATCH=$k MXX_GQ_BATCH=$k timeout 200 python tools/preimage_prof.py > $O/preimage_k$k.log 2>&1; echo "preimage k=$k rc=$? t=$(( $(date +%s) - t0 ))"
done
for o in 0 1; do
  MXX_MATMUL_COLS_FAST=$o timeout 300 python tools/md_lanes_ab.py > $O/md_order_$o.log 2>&1; echo "md order=$o rc=$? t=$(( $(date +%s) - t0 ))"
done
timeout 500 python -m pytest tests/test_samplers.py tests/test_configs_gpu.py tests/test_gpu_parity.py -m gpu -q --durations=8 > $O/pytest_subset.log 2>&1; echo "pytest rc=$? t=$(( $(date +%s) - t0 ))"
tail -15 $O/pytest_subset.log
for k in 1 8; do
  MXX_SAMPLER_BATCH=$k MXX_GQ_BATCH=$k timeout 240 ncu --metrics smsp__thread_inst_executed_per_inst_executed.ratio,smsp__inst_executed.sum,gpu__time_duration.sum \
    --clock-control none -k regex:"gauss_samp_gq|sample_gauss_batch|sample_small" -c 8 --csv --log-file $O/ncu_lanes_k$k.csv python tools/preimage_prof.py > $O/ncu_k$k.log 2>&1
  echo "ncu k=$k rc=$? t=$(( $(date +%s) - t0 ))"
done
grep -h "total\|gauss_samp\|p2 =" $O/preimage_k*.log | head -60
cat $O/md_order_0.log $O/md_order_1.log
