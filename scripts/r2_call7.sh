#!/bin/bash
# inline sun shadow rays: parity, then the headline frame and one rank's step of the 8-GPU frame with and without them
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests/test_render_gpu.py tests/test_vs_reference_gpu.py tests/test_zz_no_alloc.py -m gpu -q -x ) > gpurun_out/pytest_inline.log 2>&1; tail -5 gpurun_out/pytest_inline.log
run() {
  local name=$1; shift
  local envs=(); while [[ "$1" == *=* ]]; do envs+=("$1"); shift; done
  env "${envs[@]}" timeout 300 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-gpu-baseline "$@" > gpurun_out/sweep_$name.json 2> gpurun_out/sweep_$name.err
  python - "$name" <<PY
import json,sys
try:
    d=json.loads(open('gpurun_out/sweep_%s.json'%sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], round(d['value'],1), 'Mrays/s', round(d['ms_per_step'],2), 'ms; serial', round(d['ms_per_step_serial_passes'],2), 'ms/launch', round(d['roofline']['ms_per_launch'],3), 'launches/frame', d['gpu_launches']/4, 'nodes/ray', round(d['nodes_per_ray'],2), 'e2e', round(d['e2e']['value'],1), 'build', round(d['lbvh_build_ms'],3))
except Exception as e:
    print(sys.argv[1], 'FAILED', e); print(open('gpurun_out/sweep_%s.err'%sys.argv[1]).read()[-1500:])
PY
}
run inline
run two_launch RT_NO_INLINE_SUN=1
run inline_1080 --width 1920 --height 1080
{
python scripts/slice_step_timing.py inline 8
RT_NO_INLINE_SUN=1 python scripts/slice_step_timing.py two_launch 8
RT_MIN_PASSES=2 python scripts/slice_step_timing.py inline_two_passes 8
python scripts/slice_step_timing.py inline_n1 1
python scripts/slice_step_timing.py inline_n4 4
} 2>&1 | grep -v Warning | tee gpurun_out/slice_step_inline.txt
