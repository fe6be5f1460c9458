#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_mufreq.py -m gpu -x -q > gpurun_out/j23_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/j23_pytest.log; tail -3 gpurun_out/j23_pytest.log
timeout 900 python bench.py --workload config4 --steps 20 --warmup 5 > gpurun_out/j23_config4.json 2> gpurun_out/j23_config4.err; echo "config4 rc=$?"; tail -2 gpurun_out/j23_config4.err; python -c "
import json; d=json.loads(open('gpurun_out/j23_config4.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['value'], d['roofline'], d['e2e'])"
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:mufreq -c 12 --csv --log-file gpurun_out/j23_launches_config4.csv python bench.py --workload config4 --steps 3 --warmup 3 > gpurun_out/j23_ncu.log 2>&1
grep -v "^==" gpurun_out/j23_launches_config4.csv | cut -d, -f5,13- | grep duration | head -6
