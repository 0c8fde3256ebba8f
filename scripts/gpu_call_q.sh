#!/bin/bash
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests/test_render_gpu.py -m gpu -x -q ) > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_1m.json 2> gpurun_out/bench_1m.err
python -c "
import json
d=json.loads([l for l in open('gpurun_out/bench_1m.json').read().splitlines() if l.startswith('{')][-1])
print(round(d['value'],1), d['ms_steps_rank0'], d['e2e']['value'], d['clocks'], d['cpu_baseline']['value'], d['gpu_baseline'].get('reference_mrays_s'), d['gpu_baseline'].get('b200_mrays_s'))"
