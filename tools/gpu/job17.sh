#!/bin/bash
mkdir -p gpurun_out
SQZ_COLWARP_NW=2 timeout 600 ncu --set full --clock-control none --import-source on -k regex:baf_grid_colwarp -s 1 -c 1 -o gpurun_out/j17_nw2 -f python tools/quick_c3.py 1 > gpurun_out/j17.log 2>&1
tail -3 gpurun_out/j17.log
