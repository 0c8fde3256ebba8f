#!/bin/bash
# round 2, third GPU call (1 GPU): interleaved primitive records + 2048-key sort tiles + TMA variant (tests, build A/B),
# 1/8-slice timing, the full default bench line, the reference arm, and config 5's size on one GPU with the reference's CUDA build beside it
mkdir -p gpurun_out
L=$PWD/cuda_ray_tracer_b200
( time timeout 900 python -m pytest tests/test_build_gpu.py tests/test_render_gpu.py -m gpu -q -x ) > gpurun_out/pytest_build_render.log 2>&1; tail -3 gpurun_out/pytest_build_render.log
( RT_LIB=$L/libb200rt_tma.so timeout 600 python -m pytest tests/test_build_gpu.py -m gpu -q -x ) > gpurun_out/pytest_build_tma.log 2>&1; tail -2 gpurun_out/pytest_build_tma.log
for v in "" _s16 _tma; do RT_LIB=$L/libb200rt$v.so timeout 300 python scripts/build_ab.py "lib$v"; done 2>&1 | tee gpurun_out/build_ab.txt
timeout 300 python scripts/rank_slice_timing.py 2>&1 | tee gpurun_out/rank_slice.txt
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_1m.json 2> gpurun_out/bench_1m.err; tail -c 600 gpurun_out/bench_1m.err
timeout 900 python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; tail -c 600 gpurun_out/bench_ref.err
timeout 1500 python bench.py --tris 10000000 --spp 64 --steps 3 --warmup 3 --cpu-seconds 8 > gpurun_out/bench_10m.json 2> gpurun_out/bench_10m.err; tail -c 600 gpurun_out/bench_10m.err
python - <<PY
import json
for f in ("bench_1m", "bench_ref", "bench_10m"):
    try:
        d = json.loads(open("gpurun_out/%s.json" % f).read().strip().splitlines()[-1])
        print(f, round(d["value"], 1), "Mrays/s", d.get("ms_per_step"), "e2e", d["e2e"]["value"], "build", d.get("lbvh_build_ms"), "roofline", (d.get("roofline") or {}).get("frac"), "gpu_baseline", d.get("gpu_baseline"), "cpu", (d.get("cpu_baseline") or {}).get("value"))
    except Exception as e:
        print(f, "FAILED", e)
PY
