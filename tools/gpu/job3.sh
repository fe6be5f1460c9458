#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/j3_c3.txt
for v in 0 1 2; do SQZ_CW_VARIANT=$v timeout 300 python tools/quick_c3.py >> gpurun_out/j3_c3.txt 2>&1; done
SQZ_CW_MAXCHUNKS=64 timeout 300 python tools/quick_c3.py >> gpurun_out/j3_c3.txt 2>&1
SQZ_CW_MAXCHUNKS=64 SQZ_CW_VARIANT=1 timeout 300 python tools/quick_c3.py >> gpurun_out/j3_c3.txt 2>&1
SD_SCALE=10 timeout 300 python tools/quick_c3.py >> gpurun_out/j3_c3.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_fullsize.py tests/test_gpu_baf.py -m gpu -x -q > gpurun_out/j3_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/j3_pytest.log
(echo "== microbench"; tools/microbench; echo "== microbench2"; tools/microbench2; echo "== microbench3"; tools/microbench3) > gpurun_out/j3_microbench.txt 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:baf_grid_colwarp -s 2 -c 1 -o gpurun_out/j3_colwarp -f python tools/quick_c3.py > gpurun_out/j3_ncu_full.log 2>&1
cat gpurun_out/j3_c3.txt; tail -3 gpurun_out/j3_pytest.log
