#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/shard_balance.py 8 4 > gpurun_out/j9_balance.txt 2>&1; cat gpurun_out/j9_balance.txt
