#!/bin/bash
mkdir -p gpurun_out
SQZ_COLWARP_WS=2 timeout 600 ncu --set full --clock-control none --import-source on -k regex:baf_grid_colwarp_ws -s 1 -c 1 -o gpurun_out/j29_ws -f python tools/quick_c3.py 1 > gpurun_out/j29.log 2>&1
tail -2 gpurun_out/j29.log
