#!/bin/bash
mkdir -p gpurun_out
timeout 120 tools/mb4 > gpurun_out/j21_microbench4.txt 2>&1; cat gpurun_out/j21_microbench4.txt
