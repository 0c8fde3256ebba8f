#!/bin/bash
# build chain with programmatic dependent launch: parity of the build tests (with and without), determinism, then the A/B timing
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_build_gpu.py tests/test_config5_10m_gpu.py tests/test_determinism_gpu.py -x -q -m gpu > gpurun_out/pdl_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pdl_pytest.log
tail -4 gpurun_out/pdl_pytest.log
RT_NO_PDL=1 timeout 900 python -m pytest tests/test_build_gpu.py -x -q -m gpu > gpurun_out/pdl_pytest_off.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pdl_pytest_off.log
tail -3 gpurun_out/pdl_pytest_off.log
timeout 600 python scripts/build_ab.py HEAD > gpurun_out/pdl_ab.log 2>&1; grep -E "default|no pdl" gpurun_out/pdl_ab.log
timeout 300 python bench.py --steps 4 --warmup 3 > gpurun_out/pdl_bench.json 2> gpurun_out/pdl_bench.err; python -c "
import json; d=json.loads(open('gpurun_out/pdl_bench.json').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['e2e']['value'], d.get('lbvh_build_ms'))"
