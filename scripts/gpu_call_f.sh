#!/bin/bash
mkdir -p gpurun_out
( RT_LIB=$PWD/cuda_ray_tracer_b200/libb200rt_pair.so timeout 900 python -m pytest tests/test_render_gpu.py tests/test_vs_reference_gpu.py -m gpu -x -q ) > gpurun_out/pytest_pair.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_pair.log
tail -4 gpurun_out/pytest_pair.log
VARIANTS="${VARIANTS}" bash scripts/sweep_variants.sh 2>&1 | tee gpurun_out/sweep.txt
