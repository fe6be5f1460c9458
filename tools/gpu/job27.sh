#!/bin/bash
# final single-GPU check: suite, smoke, both bench arms, every other workload
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/j27_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/j27_pytest.log; tail -3 gpurun_out/j27_pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/j27_smoke.log 2>&1; tail -2 gpurun_out/j27_smoke.log
timeout 900 python bench.py > gpurun_out/j27_bench.json 2> gpurun_out/j27_bench.err; echo "bench rc=$?"; python tools/brief.py < gpurun_out/j27_bench.json; tail -2 gpurun_out/j27_bench.err
timeout 900 python bench.py --impl reference > gpurun_out/j27_bench_ref.json 2> gpurun_out/j27_bench_ref.err; echo "ref rc=$?"; head -c 600 gpurun_out/j27_bench_ref.json; echo
for w in config2 config4 config5 mufreq_grid segstats; do
  timeout 900 python bench.py --workload $w --steps 10 --warmup 3 > gpurun_out/j27_$w.json 2> gpurun_out/j27_$w.err; echo "$w rc=$?"
  python -c "
import json,sys
d=json.loads(open('gpurun_out/j27_$w.json').read().strip().splitlines()[-1])
print('  ', d['metric'][:50], '| ms', round(d['ms_per_step'],4), '| value %.4g' % d['value'], d['unit'], '| frac', (d.get('roofline') or {}).get('frac'), '| e2e ms', (d.get('e2e') or {}).get('ms_per_step'))"
done
