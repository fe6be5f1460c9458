#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_segstats.py tests/test_gpu_mufreq.py -m gpu -x -q > gpurun_out/j24_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/j24_pytest.log; tail -3 gpurun_out/j24_pytest.log
timeout 900 python bench.py --workload segstats --steps 20 --warmup 5 > gpurun_out/j24_segstats.json 2> gpurun_out/j24_segstats.err; echo "segstats rc=$?"; tail -2 gpurun_out/j24_segstats.err; python -c "
import json; d=json.loads(open('gpurun_out/j24_segstats.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['value'], d['roofline']['frac'])"
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:segstat -c 8 --csv --log-file gpurun_out/j24_launches_segstats.csv python bench.py --workload segstats --steps 3 --warmup 3 > gpurun_out/j24_ncu.log 2>&1
grep -v "^==" gpurun_out/j24_launches_segstats.csv | cut -d, -f5,13- | grep duration | head -4
