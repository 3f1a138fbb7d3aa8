#!/bin/bash
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out/lab6
for b in lab/mmr_*; do echo -n "$(basename $b) "; timeout 60 $b; done 2>&1 | tee gpurun_out/lab6/mmrows.txt
