#!/bin/bash
show() { python -c "import json,sys; d=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith('{')][-1]); import statistics as S; m=d['ms_steps_rank0']['device']; print(sys.argv[1], round(d['value'],1), 'median', round(S.median(m),3), 'serial', round(d['ms_per_step_serial_passes'],3), 'ms/launch', round(d['roofline']['ms_per_launch'],3))" "$1"; }
A="--steps 4 --warmup 2 --no-cpu-baseline --no-gpu-baseline"
python bench.py $A | show "default          "
RT_L2_PERSIST_NODES=1 python bench.py $A | show "L2 window on nodes"
