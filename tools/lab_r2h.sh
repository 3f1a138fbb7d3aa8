#!/bin/bash
# Round-2 closing records: whole GPU suite, the default bench line (N = 1), the ncu launch list of a short bench run, and one
# ncu --set full capture of the sampler kernels of the C5 preimage (exported to CSV on the box).
cd "${GRAFT_REPO_ROOT:-.}"
O=gpurun_out/final
mkdir -p $O
t0=$(date +%s)
timeout 600 python -m pytest tests -m gpu -q -x > $O/pytest_gpu_all.log 2>&1; echo "pytest rc=$? t=$(( $(date +%s) - t0 ))"; tail -2 $O/pytest_gpu_all.log
timeout 900 python bench.py > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$? t=$(( $(date +%s) - t0 ))"; tail -c 300 $O/bench_n1.json
timeout 500 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file $O/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu > $O/ncu_bench.log 2>&1
echo "ncu launch list rc=$? t=$(( $(date +%s) - t0 ))"; wc -l $O/launches.csv
timeout 300 ncu --set full --clock-control none --import-source on -k regex:"gauss_samp_gq_kernel|sample_gauss_batch_kernel" -s 2 -c 2 -f -o /tmp/samplers python tools/preimage_prof.py > $O/ncu_samplers.log 2>&1
echo "ncu samplers rc=$? t=$(( $(date +%s) - t0 ))"
ncu -i /tmp/samplers.ncu-rep --page raw --csv > $O/samplers_raw.csv 2>/dev/null
ncu -i /tmp/samplers.ncu-rep --page source --csv > $O/samplers_source.csv 2>/dev/null
ls -la $O
