#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/j28.txt
for ws in 0 2; do
SQZ_COLWARP_WS=$ws timeout 120 python tools/quick_c3.py >> gpurun_out/j28.txt 2>&1 || echo "ws=$ws rc=$?" >> gpurun_out/j28.txt
done
SQZ_COLWARP_WS=2 SD_SCALE=0.1 timeout 120 python tools/quick_c3.py >> gpurun_out/j28.txt 2>&1
SQZ_COLWARP_WS=2 timeout 900 python -m pytest tests/test_gpu_fullsize.py tests/test_gpu_baf.py -m gpu -x -q > gpurun_out/j28_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/j28_pytest.log
cut -c1-330 gpurun_out/j28.txt; tail -3 gpurun_out/j28_pytest.log
