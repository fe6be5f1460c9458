#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/j16.txt
SQZ_COLWARP_NW=1 timeout 300 python tools/quick_c3.py >> gpurun_out/j16.txt 2>&1
SQZ_COLWARP_NW=2 timeout 300 python tools/quick_c3.py >> gpurun_out/j16.txt 2>&1
SQZ_COLWARP_NW=26 timeout 300 python tools/quick_c3.py >> gpurun_out/j16.txt 2>&1
SQZ_COLWARP_NW=1 SD_SCALE=0.1 timeout 300 python tools/quick_c3.py >> gpurun_out/j16.txt 2>&1
WORKLOAD=c2 timeout 300 python tools/quick_c3.py 20 >> gpurun_out/j16.txt 2>&1
timeout 1200 python -m pytest tests/test_gpu_fullsize.py tests/test_gpu_baf.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/j16_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/j16_pytest.log
cut -c1-330 gpurun_out/j16.txt; tail -3 gpurun_out/j16_pytest.log
