#!/bin/bash
mkdir -p gpurun_out
for gb in 3 6 12 18; do
  RT_POOL_BYTES=$((gb << 30)) python bench.py --steps 4 --warmup 2 --no-cpu-baseline --no-gpu-baseline 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('pool ${gb} GB:', round(d['value'],1), 'Mrays/s', round(d['ms_per_step'],2), 'ms; serial', round(d['ms_per_step_serial_passes'],2), 'ms; e2e', round(d['e2e']['value'],1), 'launches', d['gpu_launches'])"
done | tee gpurun_out/pool_sweep.txt
