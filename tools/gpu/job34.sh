#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/j34.txt
for c in 96 107 116 128 140 151; do
  echo "== chunks cap $c" >> gpurun_out/j34.txt
  SQZ_CW_MAXCHUNKS=$c timeout 120 python tools/quick_c3.py >> gpurun_out/j34.txt 2>&1
  SQZ_CW_MAXCHUNKS=$c SKIP_GROUPS=1 timeout 300 python tools/shard_balance.py 8 >> gpurun_out/j34.txt 2>&1
done
cut -c1-200 gpurun_out/j34.txt
