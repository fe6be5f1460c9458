#!/bin/bash
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:segstat_kde -s 3 -c 1 -o gpurun_out/j25_kde -f python bench.py --workload segstats --steps 2 --warmup 3 > gpurun_out/j25_b.log 2>&1
echo done
