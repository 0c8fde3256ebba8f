#!/bin/bash
# rt_trace_static: alone in a fresh process, then after the whole GPU suite (pytest -x order, collected last)
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x ) > gpurun_out/pytest_noalloc_alone.log 2>&1; tail -3 gpurun_out/pytest_noalloc_alone.log
( time timeout 1700 python -m pytest tests -m gpu -q -x -rs ) > gpurun_out/pytest_gpu_full.log 2>&1; tail -8 gpurun_out/pytest_gpu_full.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
