#!/bin/bash
# round 2, first GPU call: pixel-major sample ids + packet kernel — numbers first, then the whole GPU suite
mkdir -p gpurun_out
run() {  # name, env..., lib suffix
  local name=$1; shift
  env "$@" timeout 300 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-gpu-baseline > gpurun_out/sweep_$name.json 2> gpurun_out/sweep_$name.err
  python - "$name" <<PY
import json,sys
try:
    d=json.loads(open('gpurun_out/sweep_%s.json'%sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], round(d['value'],1), 'Mrays/s', round(d['ms_per_step'],2), 'ms; serial', round(d['ms_per_step_serial_passes'],2), 'ms/launch', round(d['roofline']['ms_per_launch'],3), 'nodes/ray', round(d['nodes_per_ray'],2), 'leaf/ray', round(d['leaf_tests_per_ray'],2), 'pkt lanes', d.get('packet_lanes_per_node_step'), 'e2e', round(d['e2e']['value'],1), 'build', round(d['lbvh_build_ms'],3))
except Exception as e:
    print(sys.argv[1], 'FAILED', e); print(open('gpurun_out/sweep_%s.err'%sys.argv[1]).read()[-1500:])
PY
}
L=$PWD/cuda_ray_tracer_b200
run packet RT_PACKET=1
run perray RT_PACKET=0
run packet_p8 RT_PACKET=1 RT_LIB=$L/libb200rt_p8.so
run packet_p10 RT_PACKET=1 RT_LIB=$L/libb200rt_p10.so
( time timeout 1700 python -m pytest tests -m gpu -q -x -rs ) > gpurun_out/pytest_gpu.log 2>&1
tail -15 gpurun_out/pytest_gpu.log
