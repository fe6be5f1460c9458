#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_segstats.py -m gpu -x -q > gpurun_out/j11_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/j11_pytest.log; tail -25 gpurun_out/j11_pytest.log
timeout 900 python bench.py --workload segstats --steps 5 --warmup 3 > gpurun_out/j11_segstats.json 2> gpurun_out/j11_segstats.err; echo "segstats rc=$?"; tail -3 gpurun_out/j11_segstats.err; head -c 1500 gpurun_out/j11_segstats.json
