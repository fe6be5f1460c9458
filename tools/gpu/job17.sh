#!/bin/bash
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:baf_grid_colwarp -s 1 -c 1 -o gpurun_out/j19_c3 -f python tools/quick_c3.py 1 > gpurun_out/j19.log 2>&1
tail -3 gpurun_out/j19.log
