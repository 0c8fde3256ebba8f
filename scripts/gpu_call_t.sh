#!/bin/bash
show() { python -c "import json,sys; d=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith('{')][-1]); import statistics as S; m=d['ms_steps_rank0']['device']; print(sys.argv[1], 'median', round(S.median(m),3), 'min', min(m), 'serial', round(d['ms_per_step_serial_passes'],3))" "$1"; }
export RT_LIB=$PWD/cuda_ray_tracer_b200/libb200rt_t64.so
python bench.py --steps 6 --warmup 3 --no-cpu-baseline --no-gpu-baseline | show "4K t64"
for k in 1 2; do RT_MIN_PASSES=$k python bench.py --steps 12 --warmup 3 --no-cpu-baseline --no-gpu-baseline --width 1358 --height 764 | show "1/8 t64 passes>=$k"; done
unset RT_LIB
python bench.py --steps 6 --warmup 3 --no-cpu-baseline --no-gpu-baseline | show "4K t128"
