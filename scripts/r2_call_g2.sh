#!/bin/bash
# 2-GPU check (gpurun --gpus 2): the C-ABI communicator in one process (tests) and under torchrun (bench), before the 8-GPU call
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
nvidia-smi -L
( time timeout 240 python -m pytest tests/test_multi_gpu.py -m gpu -q -x -rs --timeout 120 --timeout-method thread ) > gpurun_out/pytest_multi_g2.log 2>&1; tail -6 gpurun_out/pytest_multi_g2.log
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 \
  > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; tail -c 1500 gpurun_out/bench_n2.err
python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/bench_n2.json").read().strip().splitlines()[-1])
    print("N=2", round(d["value"], 1), "Mrays/s", round(d["ms_per_step"], 3), "ms e2e", round(d["e2e"]["value"], 1), d.get("scaling_diag"))
except Exception as e:
    print("N=2 FAILED", e)
PY
