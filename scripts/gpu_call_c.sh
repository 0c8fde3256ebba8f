#!/bin/bash
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
VARIANTS="${VARIANTS}" bash scripts/sweep_variants.sh 2>&1 | tee gpurun_out/sweep.txt
