#!/bin/bash
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
python scripts/scene_create_bench.py | tee gpurun_out/scene_create_new.json
RT_LIB=$PWD/cuda_ray_tracer_b200/libb200rt_old.so python scripts/scene_create_bench.py | tee gpurun_out/scene_create_old.json
