#!/bin/bash
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests/test_render_gpu.py tests/test_vs_reference_gpu.py tests/test_config5_10m_gpu.py -m gpu -x -q ) > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
VARIANTS="${VARIANTS}" bash scripts/sweep_variants.sh 2>&1 | tee gpurun_out/sweep.txt
python scripts/launch_fixed_cost.py | tee gpurun_out/launch_fixed_cost_drain.json
python scripts/rank_slice_timing.py 2>&1 | head -8 | tee gpurun_out/rank_slice.txt
