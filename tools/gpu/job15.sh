#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/j15.txt
timeout 300 python tools/quick_c3.py >> gpurun_out/j15.txt 2>&1
SD_SCALE=0.1 timeout 300 python tools/quick_c3.py >> gpurun_out/j15.txt 2>&1
WORKLOAD=c2 timeout 300 python tools/quick_c3.py 20 >> gpurun_out/j15.txt 2>&1
SKIP_GROUPS=1 timeout 600 python tools/shard_balance.py 8 >> gpurun_out/j15.txt 2>&1
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/j15_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/j15_pytest.log
cut -c1-300 gpurun_out/j15.txt; tail -3 gpurun_out/j15_pytest.log
