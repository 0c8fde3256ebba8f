#!/bin/bash
show() { python -c "import json,sys; d=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith('{')][-1]); import statistics as S; m=d['ms_steps_rank0']['device']; print('$1', 'median', round(S.median(m),3), 'min', min(m), 'mean', round(d['ms_per_step'],3), 'launches', d['gpu_launches'])"; }
A="--steps 16 --warmup 3 --no-cpu-baseline --no-gpu-baseline --width 1358 --height 764"
for k in 1 2 4 1 2 4; do RT_MIN_PASSES=$k python bench.py $A | show "1/8 frame, passes>=$k"; done
