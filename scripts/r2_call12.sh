#!/bin/bash
# generate_kernel: exact multiply-shift divisions + 8-bit skip-ahead tables: parity, bench, ncu instruction count
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests/test_render_gpu.py tests/test_vs_reference_gpu.py tests/test_multi_gpu.py tests/test_zz_no_alloc.py -m gpu -q -x ) > gpurun_out/pytest_gen.log 2>&1; tail -3 gpurun_out/pytest_gen.log
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-gpu-baseline > gpurun_out/sweep_gen.json 2> gpurun_out/sweep_gen.err
python - <<PY
import json
d=json.loads(open('gpurun_out/sweep_gen.json').read().strip().splitlines()[-1])
print('fastdiv + byte tables', round(d['value'],1), 'Mrays/s', round(d['ms_per_step'],2), 'ms; serial', round(d['ms_per_step_serial_passes'],2), 'e2e', round(d['e2e']['value'],1))
PY
timeout 300 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-gpu-baseline --width 1920 --height 1080 > gpurun_out/sweep_gen_1080.json 2>/dev/null
python -c "import json; d=json.loads(open('gpurun_out/sweep_gen_1080.json').read().strip().splitlines()[-1]); print('1080p', round(d['value'],1), round(d['ms_per_step'],2))"
python scripts/slice_step_timing.py fastdiv_bytes 8 2>&1 | grep -v Warning
timeout 600 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -k regex:'generate_kernel|shade_kernel|resolve_kernel' -s 5 -c 5 --csv --log-file gpurun_out/gen_inst.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-gpu-baseline > /dev/null 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open('gpurun_out/gen_inst.csv')) if len(r)>10]
h=rows[0]; ik=h.index('Kernel Name'); im=h.index('Metric Name'); iv=h.index('Metric Value')
for r in rows[1:]: print(r[ik].split('(')[0][-40:], r[im], r[iv])
PY
