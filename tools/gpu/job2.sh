#!/bin/bash
# Round-2 GPU job 2: the software-pipelined column-per-lane kernel.
mkdir -p gpurun_out
timeout 300 python tools/quick_c3.py > gpurun_out/j2_c3.txt 2>&1
SD_SCALE=10 timeout 300 python tools/quick_c3.py >> gpurun_out/j2_c3.txt 2>&1
SD_SCALE=0.1 timeout 300 python tools/quick_c3.py >> gpurun_out/j2_c3.txt 2>&1
SQZ_CW_MAXCHUNKS=64 timeout 300 python tools/quick_c3.py >> gpurun_out/j2_c3.txt 2>&1
timeout 1500 python -m pytest tests/test_gpu_fullsize.py tests/test_gpu_baf.py tests/test_gpu_dist.py tests/test_zz_no_alloc.py -m gpu -x -q -s > gpurun_out/j2_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/j2_pytest.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:baf_grid_colwarp -s 2 -c 1 -o gpurun_out/j2_colwarp -f python tools/quick_c3.py > gpurun_out/j2_ncu_full.log 2>&1
cat gpurun_out/j2_c3.txt; tail -4 gpurun_out/j2_pytest.log
