#!/bin/bash
# 1-GPU experiment: one rank's step of the 8-GPU frame under scheduling variants (scripts/slice_step_timing.py)
mkdir -p gpurun_out
{
python scripts/slice_step_timing.py default 8
RT_NO_OVERLAP=1 python scripts/slice_step_timing.py no_overlap 8
RT_MIN_PASSES=2 python scripts/slice_step_timing.py two_passes_full_grid 8
RT_MIN_PASSES=2 RT_TRAV_CTAS_PER_SM=5 python scripts/slice_step_timing.py two_passes_5ctas 8
RT_MIN_PASSES=2 RT_TRAV_CTAS_PER_SM=4 python scripts/slice_step_timing.py two_passes_4ctas 8
RT_MIN_PASSES=4 RT_TRAV_CTAS_PER_SM=4 python scripts/slice_step_timing.py four_passes_4ctas 8
RT_WIDE4=1 python scripts/slice_step_timing.py wide4_fp32 8
python scripts/slice_step_timing.py default_n1 1
RT_TRAV_CTAS_PER_SM=5 python scripts/slice_step_timing.py n1_5ctas 1
} 2>&1 | grep -v Warning | tee gpurun_out/slice_step.txt
