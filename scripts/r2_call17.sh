#!/bin/bash
# warp climb: sibling found by one shuffle + one warp OR (default) against MATCH.ANY (libb200rt_prev.so = the commit before): parity, then A/B
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 300 python -m pytest tests/test_build_gpu.py tests/test_determinism_gpu.py -x -q -m gpu --timeout 120 --timeout-method thread > gpurun_out/climb2_pytest.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/climb2_pytest.log
D=cuda_ray_tracer_b200
for lib in libb200rt.so libb200rt_prev.so; do
  BUILD_AB_MODES=default RT_LIB=$PWD/$D/$lib timeout 120 python scripts/build_ab.py $lib
done > gpurun_out/climb2_ab.log 2>&1
cat gpurun_out/climb2_ab.log
