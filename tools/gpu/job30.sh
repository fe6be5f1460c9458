#!/bin/bash
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/j30_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/j30_pytest.log; tail -4 gpurun_out/j30_pytest.log
