#!/bin/bash
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_1m.json 2> gpurun_out/bench_1m.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_1m.json').read().strip().splitlines()[-1])
print("value", d["value"], "ms", d["ms_per_step"], "serial ms", d["ms_per_step_serial_passes"], "e2e", d["e2e"]["value"], "roof", d["roofline"]["achieved"], d["roofline"]["ms_per_launch"], d["roofline"]["share_of_step"])
PY
python scripts/rank_slice_timing.py 2>&1 | head -8 | tee gpurun_out/rank_slice.txt
