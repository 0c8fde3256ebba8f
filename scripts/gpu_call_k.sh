#!/bin/bash
mkdir -p gpurun_out
for v in "" _fixed _g16; do echo "variant '$v'"; RT_LIB=$PWD/cuda_ray_tracer_b200/libb200rt$v.so python scripts/launch_fixed_cost.py | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['warm'], d['warm_fit'])"; done 2>&1 | tee gpurun_out/fixed_cost_variants.txt
VARIANTS="_fixed _g16" bash scripts/sweep_variants.sh 2>&1 | tee gpurun_out/sweep.txt
for v in "" _fixed; do RT_LIB=$PWD/cuda_ray_tracer_b200/libb200rt$v.so python scripts/rank_slice_timing.py 2>&1 | sed -n 6,7p; done | tee gpurun_out/rank_slice.txt
