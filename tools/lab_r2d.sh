#!/bin/bash
# Sampler kernels under ncu (kernel durations, executed warp instructions, active lanes per instruction) for each
# tasks-per-lane setting; the preimage wall clock varies from box to box, the kernel times do not.
cd "${GRAFT_REPO_ROOT:-.}"
O=gpurun_out/lab4
mkdir -p $O
for k in ${LAB_KS:-1 4 8}; do
  MXX_SAMPLER_BATCH=$k MXX_GQ_BATCH=$k timeout 240 ncu --metrics smsp__thread_inst_executed_per_inst_executed.ratio,smsp__inst_executed.sum,gpu__time_duration.sum,launch__registers_per_thread \
    --clock-control none -k regex:"gauss_samp_gq|sample_gauss_batch|sample_small|p1_sample" -c 9 --csv --log-file $O/ncu_lanes_k$k.csv python tools/preimage_prof.py > $O/ncu_k$k.log 2>&1
  echo "ncu k=$k rc=$?"
done
timeout 300 python -m pytest tests/test_samplers.py -m gpu -q > $O/pytest_samplers.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_samplers.log
python - <<'PY'
import csv, glob
for f in sorted(glob.glob('gpurun_out/lab4/ncu_lanes_k*.csv')):
    rows = list(csv.reader(open(f))); hdr = None; data = {}
    for r in rows:
        if len(r) > 5 and r[0] == 'ID': hdr = r; continue
        if hdr and len(r) == len(hdr):
            d = dict(zip(hdr, r)); data.setdefault((int(d['ID']), d['Kernel Name'][:46]), {})[d['Metric Name'].split('__')[1][:14]] = d['Metric Value']
    print(f)
    for k, v in sorted(data.items()):
        if k[0] >= 2: print('  ', k, v)
PY
