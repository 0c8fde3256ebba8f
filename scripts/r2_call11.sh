#!/bin/bash
# cheaper culling key (rt_key_lb): parity suite, then the headline bench and one rank's step
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests/test_render_gpu.py tests/test_vs_reference_gpu.py tests/test_config5_10m_gpu.py tests/test_zz_no_alloc.py -m gpu -q -x ) > gpurun_out/pytest_key.log 2>&1; tail -4 gpurun_out/pytest_key.log
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-gpu-baseline > gpurun_out/sweep_key.json 2> gpurun_out/sweep_key.err
python - <<PY
import json
d=json.loads(open('gpurun_out/sweep_key.json').read().strip().splitlines()[-1])
print('cheap key', round(d['value'],1), 'Mrays/s', round(d['ms_per_step'],2), 'ms; serial', round(d['ms_per_step_serial_passes'],2), 'ms/launch', round(d['roofline']['ms_per_launch'],3), 'nodes/ray', round(d['nodes_per_ray'],3), 'leaf/ray', round(d['leaf_tests_per_ray'],3), 'e2e', round(d['e2e']['value'],1))
PY
python scripts/slice_step_timing.py cheap_key 8 2>&1 | grep -v Warning
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-gpu-baseline --tris 10000000 --spp 64 > gpurun_out/sweep_key_10m.json 2> gpurun_out/sweep_key_10m.err
python - <<PY
import json
d=json.loads(open('gpurun_out/sweep_key_10m.json').read().strip().splitlines()[-1])
print('cheap key 10M', round(d['value'],1), 'Mrays/s', round(d['ms_per_step'],2), 'ms nodes/ray', round(d['nodes_per_ray'],3), 'e2e', round(d['e2e']['value'],1))
PY
