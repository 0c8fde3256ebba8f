#!/bin/bash
# round 2, second GPU call: tuning sweep of the per-ray kernel over the pixel-major queue, launch list, ncu --set full of the traversal and of the build kernels, new tests
mkdir -p gpurun_out
run() {  # name, bench args..., env via leading VAR=VAL
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
L=$PWD/cuda_ray_tracer_b200
run base
for v in c32 c128 r16 r32 r4 l16 l4 l24; do run $v RT_LIB=$L/libb200rt_$v.so; done
run wide4 --wide4
run p1080 --width 1920 --height 1080
( time timeout 900 python -m pytest tests -m gpu -q -x -rs -k "tiny_far or packet or multi or feature" ) > gpurun_out/pytest_gpu.log 2>&1
tail -8 gpurun_out/pytest_gpu.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_r02.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-gpu-baseline > gpurun_out/ncu_list.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:traverse_near -s 4 -c 4 -f -o gpurun_out/prof_traverse_r02 \
  python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-gpu-baseline > gpurun_out/ncu_full.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'morton_kernel|onesweep_kernel|bottom_up_kernel|bounds_kernel|update_tris' -s 9 -c 9 -f -o gpurun_out/prof_build_r02 \
  python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-gpu-baseline > gpurun_out/ncu_build.log 2>&1
ls -la gpurun_out/*.ncu-rep
