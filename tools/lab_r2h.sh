#!/bin/bash
# Round-2 closing records: the default bench line (N = 1), then the ncu launch list of a short bench run.
cd "${GRAFT_REPO_ROOT:-.}"
O=gpurun_out/final
mkdir -p $O
t0=$(date +%s)
timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_configs_gpu.py tests/test_zz_no_alloc.py -m gpu -q -x > $O/pytest_subset.log 2>&1; echo "pytest rc=$?"; tail -2 $O/pytest_subset.log
timeout 900 python bench.py > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$? t=$(( $(date +%s) - t0 ))"; tail -c 600 $O/bench_n1.json
timeout 500 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file $O/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu > $O/ncu_bench.log 2>&1
echo "ncu launch list rc=$? t=$(( $(date +%s) - t0 ))"; wc -l $O/launches.csv
