#!/bin/bash
# Round-2 GPU job 1: parity suite, first bench line of the reworked launch path, microbenchmarks,
# ncu launch list and one full capture of the column-per-lane kernel.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/j1_smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q -s > gpurun_out/j1_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/j1_pytest.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/j1_bench.json 2> gpurun_out/j1_bench.err; echo "bench rc=$?"
SQZ_CW_MAXCHUNKS=64 timeout 300 python tools/quick_c3.py > gpurun_out/j1_chunks64.txt 2>&1
timeout 300 python tools/quick_c3.py > gpurun_out/j1_chunks128.txt 2>&1
(echo "== microbench"; tools/microbench; echo "== microbench2"; tools/microbench2; echo "== microbench3"; tools/microbench3) > gpurun_out/j1_microbench.txt 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/j1_launches.csv python bench.py --steps 2 --warmup 3 > gpurun_out/j1_ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:baf_grid_colwarp -s 2 -c 1 -o gpurun_out/j1_colwarp -f python tools/quick_c3.py > gpurun_out/j1_ncu_full.log 2>&1
tail -3 gpurun_out/j1_pytest.log; head -c 600 gpurun_out/j1_bench.json; cat gpurun_out/j1_chunks64.txt gpurun_out/j1_chunks128.txt
