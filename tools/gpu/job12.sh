#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_segstats.py -m gpu -x -q > gpurun_out/j12_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/j12_pytest.log; tail -5 gpurun_out/j12_pytest.log
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,sm__warps_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:segstat -c 6 --csv --log-file gpurun_out/j12_launches_segstats.csv python bench.py --workload segstats --steps 3 --warmup 3 > gpurun_out/j12_ncu.log 2>&1
grep -v "^==" gpurun_out/j12_launches_segstats.csv | cut -d, -f5,13- | head -40
