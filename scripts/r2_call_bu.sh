#!/bin/bash
# climb variants of the fused builder: parity of the build tests under each switch, then the A/B timing
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
RT_BU_PEEK=1 timeout 900 python -m pytest tests/test_build_gpu.py tests/test_config5_10m_gpu.py -x -q -m gpu > gpurun_out/bu_pytest_peek.log 2>&1; echo "pytest rc=$?" >> gpurun_out/bu_pytest_peek.log
tail -3 gpurun_out/bu_pytest_peek.log
RT_BU_NO_COMPACT=1 timeout 900 python -m pytest tests/test_build_gpu.py -x -q -m gpu > gpurun_out/bu_pytest_nocompact.log 2>&1; echo "pytest rc=$?" >> gpurun_out/bu_pytest_nocompact.log
tail -3 gpurun_out/bu_pytest_nocompact.log
timeout 600 python scripts/build_ab.py HEAD > gpurun_out/bu_ab2.log 2>&1; cat gpurun_out/bu_ab2.log
