#!/bin/bash
cd "${GRAFT_REPO_ROOT:-.}"
O=gpurun_out/final3
mkdir -p $O
timeout 600 python bench.py > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$?"
python - <<PY
import json
d = json.load(open("$O/bench_n1.json"))
print(d["value"], d["e2e"]["value"], d["bgg_chain"]["n1"], d["bgg_chain_28bit"]["n1"]["ms_per_gate"], d["preimage"]["ms_per_preimage"])
PY
