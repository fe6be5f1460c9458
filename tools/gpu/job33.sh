#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/j33.txt
SQZ_COLWARP_SERIAL=1 timeout 120 python tools/quick_c3.py >> gpurun_out/j33.txt 2>&1
timeout 120 python tools/quick_c3.py >> gpurun_out/j33.txt 2>&1
SQZ_COLWARP_SERIAL=1 WORKLOAD=c2 timeout 120 python tools/quick_c3.py 20 >> gpurun_out/j33.txt 2>&1
WORKLOAD=c2 timeout 120 python tools/quick_c3.py 20 >> gpurun_out/j33.txt 2>&1
SKIP_GROUPS=1 timeout 600 python tools/shard_balance.py 8 >> gpurun_out/j33.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/j33_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/j33_pytest.log
cut -c1-330 gpurun_out/j33.txt; tail -3 gpurun_out/j33_pytest.log
