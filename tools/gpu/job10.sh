#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/j10.txt
timeout 300 python tools/quick_c3.py >> gpurun_out/j10.txt 2>&1
WORKLOAD=c2 timeout 300 python tools/quick_c3.py 20 >> gpurun_out/j10.txt 2>&1
timeout 600 python tools/shard_balance.py 8 >> gpurun_out/j10.txt 2>&1
for ch in 128 136 204; do
SQZ_CW_MAXCHUNKS=$ch timeout 300 python tools/quick_c3.py >> gpurun_out/j10.txt 2>&1
SQZ_CW_MAXCHUNKS=$ch SKIP_GROUPS=1 timeout 600 python tools/shard_balance.py 8 >> gpurun_out/j10.txt 2>&1
done
timeout 900 python -m pytest tests/test_gpu_fullsize.py tests/test_gpu_baf.py -m gpu -x -q > gpurun_out/j10_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/j10_pytest.log
cut -c1-330 gpurun_out/j10.txt; tail -3 gpurun_out/j10_pytest.log
