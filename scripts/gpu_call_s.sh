#!/bin/bash
for v in "" _nopf "" _nopf; do echo "variant '$v'"; RT_LIB=$PWD/cuda_ray_tracer_b200/libb200rt$v.so python scripts/launch_fixed_cost.py | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['warm'][:3], d['warm_fit'])"; done
show() { python -c "import json,sys; d=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith('{')][-1]); import statistics as S; m=d['ms_steps_rank0']['device']; print('$1', 'median', round(S.median(m),3), 'min', min(m), 'mean', round(d['ms_per_step'],3))"; }
A="--steps 16 --warmup 3 --no-cpu-baseline --no-gpu-baseline"
for v in "" _nopf "" _nopf; do RT_LIB=$PWD/cuda_ray_tracer_b200/libb200rt$v.so python bench.py $A --width 1358 --height 764 | show "1/8 frame '$v'"; done
for v in "" _nopf; do RT_LIB=$PWD/cuda_ray_tracer_b200/libb200rt$v.so python bench.py --steps 6 --warmup 3 --no-cpu-baseline --no-gpu-baseline | show "4K frame '$v'"; done
