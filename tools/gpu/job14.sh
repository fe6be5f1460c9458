#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_mufreq.py -m gpu -x -q > gpurun_out/j14_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/j14_pytest.log; tail -3 gpurun_out/j14_pytest.log
timeout 900 python bench.py --workload config4 --steps 20 --warmup 5 > gpurun_out/j14_config4.json 2> gpurun_out/j14_config4.err; echo "config4 rc=$?"; tail -2 gpurun_out/j14_config4.err; head -c 1800 gpurun_out/j14_config4.json
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:mufreq -c 12 --csv --log-file gpurun_out/j14_launches_config4.csv python bench.py --workload config4 --steps 3 --warmup 3 > gpurun_out/j14_ncu.log 2>&1
grep -v "^==" gpurun_out/j14_launches_config4.csv | cut -d, -f5,14- | head -30
timeout 600 python bench.py --workload mufreq_grid --steps 5 --warmup 3 > gpurun_out/j14_mufreq_grid.json 2> gpurun_out/j14_mufreq_grid.err; echo "mufreq_grid rc=$?"; head -c 900 gpurun_out/j14_mufreq_grid.json
