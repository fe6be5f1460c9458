#!/bin/bash
# config 2 (CNt 0..20, 91 x 61): split kernel against the column-per-lane kernel with narrow tiles
mkdir -p gpurun_out; rm -f gpurun_out/j5.txt
export WORKLOAD=c2
for sd in 1 0.1 10; do
  export SD_SCALE=$sd
  timeout 300 python tools/quick_c3.py 20 >> gpurun_out/j5.txt 2>&1
  for pt in 8 16; do
    for ch in 16 32 64; do
      SQZ_COLWARP=1 SQZ_PT=$pt SQZ_CHUNKS=$ch timeout 300 python tools/quick_c3.py 20 >> gpurun_out/j5.txt 2>&1
    done
  done
done
unset WORKLOAD SD_SCALE
SD_SCALE=0.1 timeout 300 python tools/quick_c3.py >> gpurun_out/j5.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_baf.py -m gpu -x -q -k "column_lane or paths" > gpurun_out/j5_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/j5_pytest.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/j5_launches_c2.csv python bench.py --steps 2 --warmup 3 --workload config2 > gpurun_out/j5_ncu_c2.log 2>&1
cat gpurun_out/j5.txt | cut -c1-260; tail -3 gpurun_out/j5_pytest.log
