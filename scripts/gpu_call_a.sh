#!/bin/bash
# round-1 GPU call: full GPU test suite, default bench, config-5-size bench on one GPU, one ncu capture of its traversal
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader > gpurun_out/gpu.txt
( time timeout 900 python -m pytest tests -m gpu -x -q --durations=15 ) > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
timeout 400 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_1m.json 2> gpurun_out/bench_1m.err
timeout 400 python bench.py --tris 10000000 --spp 64 --steps 3 --warmup 3 --no-gpu-baseline --cpu-seconds 8 > gpurun_out/bench_10m.json 2> gpurun_out/bench_10m.err
timeout 400 ncu --set full --clock-control none --import-source on -k regex:traverse_near -s 6 -c 4 -f -o gpurun_out/prof_traverse_10m \
  python bench.py --tris 10000000 --spp 64 --steps 1 --warmup 1 --no-cpu-baseline --no-gpu-baseline > gpurun_out/ncu_10m.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/bench_1m.json | cut -c1-400; cat gpurun_out/bench_10m.json | cut -c1-600
