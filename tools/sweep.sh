for pt in 32 16 8; do for th in 256 128; do for ch in 0; do
echo "PT=$pt TH=$th CH=$ch"; SQZ_PT=$pt SQZ_THREADS=$th SQZ_CHUNKS=$ch python bench.py --steps 30 --warmup 3 2>&1 | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); r=d['roofline']; print('  ms/step %.3f kernel_ms %.3f evals/s %.3e pipe_util %.3f chunks %s'%(d['ms_per_step'], r['kernel_ms'], d['value'], r['pipe_util'] or 0, d['config']['segment_chunks']))
"
done; done; done
