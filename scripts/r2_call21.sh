#!/bin/bash
# one rank's step of an 8-rank frame on one GPU at the round's last commit (profiles/slice_step_variants_r02.txt has the earlier lines)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 40 python scripts/slice_step_timing.py last_commit 8 2>&1 | grep -v Warning | tee gpurun_out/slice_last.log
