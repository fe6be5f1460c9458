#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/j6.txt
export WORKLOAD=c2 SQZ_CW_MAXCHUNKS=256
for sd in 1 0.1; do
  export SD_SCALE=$sd
  for ch in 43 49 86 98 128 172 196; do
    SQZ_COLWARP=1 SQZ_PT=8 SQZ_CHUNKS=$ch timeout 300 python tools/quick_c3.py 20 >> gpurun_out/j6.txt 2>&1
  done
  for ch in 74 98 148 196; do
    SQZ_CW_B1=1 SQZ_COLWARP=1 SQZ_PT=8 SQZ_CHUNKS=$ch timeout 300 python tools/quick_c3.py 20 >> gpurun_out/j6.txt 2>&1
  done
done
cut -c1-60,150-280 gpurun_out/j6.txt
