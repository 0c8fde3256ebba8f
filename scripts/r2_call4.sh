#!/bin/bash
# round 2, fourth GPU call (1 GPU): the reworked 4-wide step (fp32 nodes, and bfloat16-offset nodes) against the binary tree:
# parity tests, headline bench, per-launch fixed cost, 1/8-slice timing; the deep-recursion test
mkdir -p gpurun_out
L=$PWD/cuda_ray_tracer_b200
( time timeout 900 python -m pytest tests/test_build_gpu.py tests/test_render_gpu.py -m gpu -q -x -k "wide4 or deep_recursion or build or tiny_far or schedules" ) > gpurun_out/pytest_w4.log 2>&1; tail -3 gpurun_out/pytest_w4.log
( RT_LIB=$L/libb200rt_w4c.so timeout 600 python -m pytest tests/test_render_gpu.py -m gpu -q -x -k "wide4" -s ) > gpurun_out/pytest_w4c.log 2>&1; tail -4 gpurun_out/pytest_w4c.log
run() {
  local name=$1; shift
  local envs=(); while [[ "$1" == *=* ]]; do envs+=("$1"); shift; done
  env "${envs[@]}" timeout 300 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-gpu-baseline "$@" > gpurun_out/sweep_$name.json 2> gpurun_out/sweep_$name.err
  python - "$name" <<PY
import json,sys
try:
    d=json.loads(open('gpurun_out/sweep_%s.json'%sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], round(d['value'],1), 'Mrays/s', round(d['ms_per_step'],2), 'ms; serial', round(d['ms_per_step_serial_passes'],2), 'ms/launch', round(d['roofline']['ms_per_launch'],3), 'nodes/ray', round(d['nodes_per_ray'],2), 'leaf/ray', round(d['leaf_tests_per_ray'],2), 'e2e', round(d['e2e']['value'],1), 'build', round(d['lbvh_build_ms'],3))
except Exception as e:
    print(sys.argv[1], 'FAILED', e); print(open('gpurun_out/sweep_%s.err'%sys.argv[1]).read()[-1500:])
PY
}
run binary
run wide4_fp32 --wide4
run wide4_bf16 RT_LIB=$L/libb200rt_w4c.so --wide4
run wide4_bf16_1080 RT_LIB=$L/libb200rt_w4c.so --wide4 --width 1920 --height 1080
for v in "binary::" "wide4_fp32:RT_WIDE4=1:" "wide4_bf16:RT_WIDE4=1:$L/libb200rt_w4c.so"; do
  n=${v%%:*}; r=${v#*:}; w=${r%%:*}; lib=${r#*:}
  env ${w:+$w} ${lib:+RT_LIB=$lib} timeout 200 python scripts/launch_fixed_cost.py > gpurun_out/fixed_cost_$n.json 2>/dev/null
  python -c "import json; d=json.load(open('gpurun_out/fixed_cost_$n.json')); print('$n fixed cost', d['warm_fit'], d['cold_fit'])"
done
RT_WIDE4=1 RT_LIB=$L/libb200rt_w4c.so timeout 200 python scripts/rank_slice_timing.py 2>&1 | head -8 | tee gpurun_out/rank_slice_w4c.txt
