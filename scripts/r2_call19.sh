#!/bin/bash
# config 4 as named (1M soup, 1920x1080, 16 spp, rebuilt per frame) at the round's last commit
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 70 python bench.py --width 1920 --height 1080 --no-cpu-baseline --no-gpu-baseline > gpurun_out/bench_1080p.json 2> gpurun_out/bench_1080p.err
python -c "
import json; d=json.loads(open('gpurun_out/bench_1080p.json').read().strip().splitlines()[-1]); print(round(d['value'],1), d['ms_per_step'], round(d['e2e']['value'],1), d.get('lbvh_build_ms'))"
