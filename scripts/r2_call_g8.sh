#!/bin/bash
# 8-GPU call (gpurun --gpus 8): strong scaling of the headline config at N = 8, 4, 2, 1 on one box, config 5 (10M triangles,
# 4K, 64 spp) at N = 8, and the multi-GPU tests of the C ABI
mkdir -p gpurun_out
nvidia-smi -L | wc -l
run() {  # N, name, extra args
  local n=$1 name=$2; shift 2
  if [ "$n" = 1 ]; then timeout 600 python bench.py --gpus 1 "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err
  else timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) bench.py --gpus $n "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; fi
  python - "$name" <<PY
import json, sys
try:
    d = json.loads(open("gpurun_out/%s.json" % sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], "N=%d" % d["n_gpus"], round(d["value"], 1), "Mrays/s", round(d["ms_per_step"], 3), "ms; e2e", round(d["e2e"]["value"], 1), "build", round(d["lbvh_build_ms"], 3), d.get("scaling_diag"))
except Exception as e:
    print(sys.argv[1], "FAILED", e); print(open("gpurun_out/%s.err" % sys.argv[1]).read()[-1200:])
PY
}
run 8 scale_1m_n8 --steps 10 --warmup 3 --no-cpu-baseline --no-gpu-baseline
run 8 config5_10m_n8 --tris 10000000 --spp 64 --steps 4 --warmup 3 --no-cpu-baseline --no-gpu-baseline
run 4 scale_1m_n4 --steps 10 --warmup 3 --no-cpu-baseline --no-gpu-baseline
run 2 scale_1m_n2 --steps 10 --warmup 3 --no-cpu-baseline --no-gpu-baseline
run 1 scale_1m_n1 --steps 10 --warmup 3 --no-cpu-baseline --no-gpu-baseline
run 4 config5_10m_n4 --tris 10000000 --spp 64 --steps 3 --warmup 2 --no-cpu-baseline --no-gpu-baseline
( time timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -q -x -rs ) > gpurun_out/pytest_multi_g8.log 2>&1; tail -4 gpurun_out/pytest_multi_g8.log
