#!/bin/bash
# locate the test that did not finish in r2_call15 (pytest was cut at 900 s with its output still buffered): verbose, unbuffered,
# a per-test limit that dumps the stacks
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 240 python -m pytest tests/test_build_gpu.py -x -v --timeout 40 --timeout-method thread > gpurun_out/diag_build.log 2>&1; echo "build rc=$?"
grep -c PASSED gpurun_out/diag_build.log; grep -v PASSED gpurun_out/diag_build.log | tail -15
timeout 150 python -m pytest tests/test_config5_10m_gpu.py -x -v --timeout 100 --timeout-method thread > gpurun_out/diag_c5.log 2>&1; echo "config5 rc=$?"
grep -v PASSED gpurun_out/diag_c5.log | tail -8
timeout 200 python -m pytest tests/test_determinism_gpu.py -x -v --timeout 150 --timeout-method thread > gpurun_out/diag_det.log 2>&1; echo "determinism rc=$?"
grep -v PASSED gpurun_out/diag_det.log | tail -12
