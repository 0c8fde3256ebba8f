#!/bin/bash
# GPU call: full GPU test suite (-s for the printed timings), default bench
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -m gpu -x -q -s --durations=10 ) > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
timeout 400 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_1m.json 2> gpurun_out/bench_1m.err
grep -E "build ms|passed|failed|exit" gpurun_out/pytest_gpu.log | tail; cut -c1-300 gpurun_out/bench_1m.json; grep -o '"lbvh_build_ms[a-z_]*": [0-9.]*' gpurun_out/bench_1m.json
