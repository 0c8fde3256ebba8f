#!/bin/bash
# ncu --set full of the non-traversal kernels of the frame: generate, shade (first / second visit), resolve
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'generate_kernel|shade_kernel|resolve_kernel' -s 5 -c 5 -f -o gpurun_out/prof_other_r02 \
  python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-gpu-baseline > gpurun_out/ncu_other.log 2>&1
ls -la gpurun_out/prof_other_r02.ncu-rep
