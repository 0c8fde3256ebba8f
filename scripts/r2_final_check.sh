#!/bin/bash
# round-2 final check on one GPU: whole GPU suite (-x, the order the driver uses), smoke, default bench line, reference arm
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
( time RT_REQUIRE_REF=1 timeout 600 python -m pytest tests -m gpu -q -x -rs --timeout 200 --timeout-method thread ) > gpurun_out/pytest_gpu_final.log 2>&1; tail -7 gpurun_out/pytest_gpu_final.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 900 python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err; tail -c 300 gpurun_out/bench_final.err
timeout 900 python bench.py --impl reference > gpurun_out/bench_ref_final.json 2> gpurun_out/bench_ref_final.err; tail -c 300 gpurun_out/bench_ref_final.err
python - <<PY
import json
for f in ("bench_final", "bench_ref_final"):
    try:
        d = json.loads(open("gpurun_out/%s.json" % f).read().strip().splitlines()[-1])
        r = d.get("roofline") or {}
        print(f, round(d["value"], 1), "Mrays/s", d.get("ms_per_step"), "e2e", round(d["e2e"]["value"], 1), "build", d.get("lbvh_build_ms"), "| roofline", r.get("bound"), r.get("frac"), r.get("peak"), "| launches", d.get("gpu_launches"), "| clocks", d.get("clocks"), "| gpu_baseline", (d.get("gpu_baseline") or {}).get("reference_mrays_s"), "| cpu", (d.get("cpu_baseline") or {}).get("value"), d.get("reference_kind"))
    except Exception as e:
        print(f, "FAILED", e)
PY
