#!/bin/bash
# per-kernel launch list of the build chain at the round's last commit (short: the round's last GPU seconds)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 80 ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches_build.csv python scripts/build_launch_list.py > gpurun_out/launches_build.log 2>&1
tail -3 gpurun_out/launches_build.log
