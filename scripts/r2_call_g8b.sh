#!/bin/bash
# 8-GPU check of the new tile shape (256x8): headline frame at N = 8 and N = 1 on the same box, multi-GPU C-ABI tests
mkdir -p gpurun_out
run() {
  local n=$1 name=$2; shift 2
  if [ "$n" = 1 ]; then timeout 600 python bench.py --gpus 1 "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err
  else timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) bench.py --gpus $n "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; fi
  python - "$name" <<PY
import json, sys
try:
    d = json.loads(open("gpurun_out/%s.json" % sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], "N=%d" % d["n_gpus"], round(d["value"], 1), "Mrays/s", round(d["ms_per_step"], 3), "ms; e2e", round(d["e2e"]["value"], 1), "build", round(d["lbvh_build_ms"], 3), d.get("scaling_diag"), d["config"]["parallelism"][:60])
except Exception as e:
    print(sys.argv[1], "FAILED", e); print(open("gpurun_out/%s.err" % sys.argv[1]).read()[-1200:])
PY
}
run 8 scale2_1m_n8 --steps 10 --warmup 3 --no-cpu-baseline --no-gpu-baseline
run 1 scale2_1m_n1 --steps 10 --warmup 3 --no-cpu-baseline --no-gpu-baseline
run 4 scale2_1m_n4 --steps 10 --warmup 3 --no-cpu-baseline --no-gpu-baseline
( timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -q -x -rs ) 2>&1 | tail -2
