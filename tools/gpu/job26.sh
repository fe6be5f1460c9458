#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_mufreq.py -m gpu -x -q > gpurun_out/j26_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/j26_pytest.log; tail -25 gpurun_out/j26_pytest.log
