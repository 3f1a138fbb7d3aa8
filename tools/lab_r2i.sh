#!/bin/bash
cd "${GRAFT_REPO_ROOT:-.}"
O=gpurun_out/final2
mkdir -p $O
nvidia-smi --query-gpu=clocks.sm,clocks.max.sm,power.draw,clocks_throttle_reasons.active --format=csv > $O/smi.txt
for i in 1 2; do
  timeout 300 python bench.py --no-cpu > $O/bench_$i.json 2> $O/bench_$i.err; echo "bench $i rc=$?"
  python - <<PY
import json
d = json.load(open("$O/bench_$i.json"))
print(d["value"], d["e2e"]["value"], d["bgg_chain"]["n1"]["ms_per_gate_runs"], d["bgg_chain_28bit"]["n1"]["ms_per_gate_runs"], d["preimage"]["ms_per_preimage"])
PY
done
MXX_MATMUL_ROWS4=1 timeout 100 python tools/md_lanes_ab.py 2>&1 | grep "C4 gate\|C3-b"
