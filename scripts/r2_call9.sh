#!/bin/bash
# pool-size curve (RT_POOL_BYTES per pool, two pools), then the launch list and ncu --set full captures at HEAD
mkdir -p gpurun_out
for gb in 1 2 4 8 16 32; do
  RT_POOL_BYTES=$((gb << 30)) timeout 300 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-gpu-baseline > gpurun_out/pool_$gb.json 2> gpurun_out/pool_$gb.err
  python - $gb <<PY
import json,sys
try:
    d=json.loads(open('gpurun_out/pool_%s.json'%sys.argv[1]).read().strip().splitlines()[-1])
    print('pool %2s GB x 2: %.1f Mrays/s  %.2f ms/frame  e2e %.1f  launches/frame %.0f  traverse launches/frame %.0f' % (sys.argv[1], d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches']/4, d['roofline']['bytes_per_launch'] and (d['rays_per_step']*0+ d['gpu_launches']/4)))
except Exception as e:
    print('pool', sys.argv[1], 'FAILED', e); print(open('gpurun_out/pool_%s.err'%sys.argv[1]).read()[-600:])
PY
done 2>&1 | tee gpurun_out/pool_curve.txt
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_head.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-gpu-baseline > gpurun_out/ncu_list.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:traverse_near -s 4 -c 4 -f -o gpurun_out/prof_traverse_head \
  python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-gpu-baseline > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out/*.ncu-rep
