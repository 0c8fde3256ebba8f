#!/bin/bash
# one GPU call that checks a round (usage: gpurun --timeout 2400 -- bash scripts/gpu_round_check.sh): full GPU suite, smoke, default bench (+ reference arm), config-5-size bench, ncu launch list and
# one --set full capture of the traversal kernel of the bench command
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader > gpurun_out/gpu.txt
( time timeout 1200 python -m pytest tests -m gpu -x -q -s ) > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_1m.json 2> gpurun_out/bench_1m.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
timeout 600 python bench.py --tris 10000000 --spp 64 --steps 3 --warmup 3 --no-gpu-baseline --cpu-seconds 8 > gpurun_out/bench_10m.json 2> gpurun_out/bench_10m.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-gpu-baseline > gpurun_out/ncu_list.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:traverse_near -s 4 -c 4 -f -o gpurun_out/prof_traverse_r01 \
  python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-gpu-baseline > gpurun_out/ncu_full.log 2>&1
grep -E "passed|failed|exit|build ms" gpurun_out/pytest_gpu.log | tail -8; tail -2 gpurun_out/smoke.log
for f in bench_1m bench_10m; do python - <<PY
import json
d=json.loads(open('gpurun_out/$f.json').read().strip().splitlines()[-1])
print("$f", round(d["value"],1), "Mrays/s", round(d["ms_per_step"],2), "ms; serial", round(d["ms_per_step_serial_passes"],2), "; e2e", round(d["e2e"]["value"],1), "; build", round(d["lbvh_build_ms"],3), "; roofline", round(d["roofline"]["achieved"]), round(d["roofline"]["frac"],3), d["roofline"].get("frac_of_l2_gather"))
PY
done
