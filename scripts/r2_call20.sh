#!/bin/bash
# build times at the round's last commit (default switches only)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
BUILD_AB_MODES=default,global timeout 60 python scripts/build_ab.py HEAD > gpurun_out/build_final.log 2>&1; cat gpurun_out/build_final.log
