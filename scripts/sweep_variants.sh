# usage: VARIANTS="_a _b" bash scripts/sweep_variants.sh   (libb200rt$v.so built with make LIB=... EXTRA=...)
# prints: variant, Mrays/s, ms per frame (two passes in flight), ms per frame (one pass at a time), ms per traverse launch, nodes per ray
for v in "" $VARIANTS; do
  RT_LIB=$PWD/cuda_ray_tracer_b200/libb200rt$v.so python bench.py --steps 4 --warmup 2 --no-cpu-baseline --no-gpu-baseline 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$v', round(d['value'],1), round(d['ms_per_step'],2), round(d['ms_per_step_serial_passes'],2), round(d['roofline']['ms_per_launch'],3), round(d['nodes_per_ray'],2))"
done
