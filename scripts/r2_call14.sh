#!/bin/bash
mkdir -p gpurun_out
python scripts/tile_shape_scan.py 8 2>&1 | grep -v Warning | tee gpurun_out/tile_shape_scan.txt
