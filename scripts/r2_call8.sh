#!/bin/bash
# the reference's CUDA build on the largest soup its parser can load correctly (5.5M triangles: its float-typed vertex index,
# parse.cpp:175-180, breaks at 2^24 vertices = 5 592 405 triangles), 4K, aa 32, against this library on the same file
mkdir -p gpurun_out
timeout 2000 python scripts/bench_vs_gpu_reference.py --tris 5500000 --w 3840 --h 2160 --aa 32 --repeat 2 --fast-only --out gpurun_out/vs_gpu_reference_5p5m.json 2> gpurun_out/vs5m.err | cut -c1-1800
tail -c 400 gpurun_out/vs5m.err
