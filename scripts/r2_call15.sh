#!/bin/bash
# sort changes (hybrid histogram in morton_kernel, windowed look-back in onesweep): parity, A/B against variant libraries;
# then the round's final captures at HEAD: launch list, ncu --set full of the traversal and of the build kernels
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_build_gpu.py tests/test_config5_10m_gpu.py tests/test_determinism_gpu.py -x -q -m gpu > gpurun_out/sort_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/sort_pytest.log
tail -3 gpurun_out/sort_pytest.log
D=cuda_ray_tracer_b200
for lib in libb200rt.so libb200rt_lb1.so libb200rt_old.so; do
  BUILD_AB_MODES=default RT_LIB=$PWD/$D/$lib timeout 300 python scripts/build_ab.py $lib
done > gpurun_out/sort_ab.log 2>&1
cat gpurun_out/sort_ab.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_head.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-gpu-baseline > gpurun_out/ncu_list.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:traverse_near -s 4 -c 4 -f -o gpurun_out/prof_traverse_head \
  python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-gpu-baseline > gpurun_out/ncu_full.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'morton_kernel|onesweep_kernel|bottom_up|bounds_kernel|update_tris' -s 8 -c 8 -f -o gpurun_out/prof_build_head \
  python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-gpu-baseline > gpurun_out/ncu_build.log 2>&1
ls -la gpurun_out/*.ncu-rep
