#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/j18.txt
timeout 300 python tools/quick_c3.py >> gpurun_out/j18.txt 2>&1
SD_SCALE=0.1 timeout 300 python tools/quick_c3.py >> gpurun_out/j18.txt 2>&1
SD_SCALE=10 timeout 300 python tools/quick_c3.py >> gpurun_out/j18.txt 2>&1
WORKLOAD=c2 timeout 300 python tools/quick_c3.py 20 >> gpurun_out/j18.txt 2>&1
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/j18_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/j18_pytest.log
cut -c1-330 gpurun_out/j18.txt; tail -3 gpurun_out/j18_pytest.log
