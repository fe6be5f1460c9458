#!/bin/bash
mkdir -p gpurun_out
timeout 900 python bench.py --workload config4 --steps 10 --warmup 3 > gpurun_out/j32_config4.json 2> gpurun_out/j32_config4.err; echo "config4 rc=$?"; tail -3 gpurun_out/j32_config4.err; python -c "
import json; d=json.loads(open('gpurun_out/j32_config4.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['e2e'], d['e2e_pinned_zero_copy'])"
