#!/bin/bash
# guided chunk sizes: 128-ray tickets, 32-ray tickets for the last 4 rays per resident lane (default); variants
mkdir -p gpurun_out
L=$PWD/cuda_ray_tracer_b200
for v in "" _t8 _t16c16 ""; do
  RT_LIB=$L/libb200rt$v.so timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-gpu-baseline > gpurun_out/sweep_guided$v.json 2>/dev/null
  python -c "import json; d=json.loads(open('gpurun_out/sweep_guided$v.json').read().strip().splitlines()[-1]); print('guided$v', round(d['value'],1), round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value'],1))"
  RT_LIB=$L/libb200rt$v.so python scripts/slice_step_timing.py "guided$v" 8 2>&1 | grep -v Warning
done
( timeout 900 python -m pytest tests/test_render_gpu.py -m gpu -q -x ) 2>&1 | tail -2
