#!/bin/bash
mkdir -p gpurun_out
show() { python -c "import json,sys; d=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith('{')][-1]); print('$1', round(d['value'],1), d['ms_steps_rank0'])"; }
A="--steps 8 --warmup 3 --no-cpu-baseline --no-gpu-baseline"
python bench.py $A | show "16GB pools"
RT_POOL_BYTES=$((6<<30)) python bench.py $A | show "6GB pools "
python bench.py $A | show "16GB pools"
RT_POOL_BYTES=$((6<<30)) python bench.py $A | show "6GB pools "
