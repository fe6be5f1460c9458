#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/j35.txt
timeout 120 python tools/quick_c3.py >> gpurun_out/j35.txt 2>&1
SD_SCALE=0.1 timeout 120 python tools/quick_c3.py >> gpurun_out/j35.txt 2>&1
WORKLOAD=c2 timeout 120 python tools/quick_c3.py 20 >> gpurun_out/j35.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_fullsize.py tests/test_gpu_baf.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/j35_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/j35_pytest.log
cut -c1-200 gpurun_out/j35.txt; tail -3 gpurun_out/j35_pytest.log
