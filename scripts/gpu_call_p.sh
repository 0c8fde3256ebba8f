#!/bin/bash
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
show() { python -c "import json,sys; d=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith('{')][-1]); print('$1', round(d['value'],1), round(d['ms_per_step'],3), 'serial', round(d['ms_per_step_serial_passes'],3), 'e2e', round(d['e2e']['ms_per_step'],3), d['ms_steps_rank0']['device'], 'launches', d['gpu_launches'], 'roof', round(d['roofline']['achieved']), round(d['roofline']['ms_per_launch'],3))"; }
A="--steps 8 --warmup 3 --no-cpu-baseline --no-gpu-baseline"
python bench.py $A | show "4K queued     "
RT_HOST_DRIVEN=1 python bench.py $A | show "4K host-driven"
python bench.py $A --width 1358 --height 764 | show "1/8 queued     "
RT_HOST_DRIVEN=1 python bench.py $A --width 1358 --height 764 | show "1/8 host-driven"
python bench.py $A | show "4K queued     "
