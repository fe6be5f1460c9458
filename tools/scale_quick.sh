#!/bin/bash
# One GPU count of the scaling table: bash tools/scale_quick.sh N [extra]
# -> gpurun_out/scale_<workload>_n<N>.json; with "extra" also the single-process n_gpus path
# and the cohort (config 5).  tools/scale_run.sh runs the whole 1/2/4/8 series.
n=${1:-8}
port=29700
for w in config2 config3; do
  steps=30; [ $w = config3 ] && steps=5
  port=$((port+1))
  if [ $n -eq 1 ]; then
    timeout 120 python bench.py --gpus 1 --steps $steps --warmup 3 --workload $w 2>/dev/null | tail -1 > gpurun_out/scale_${w}_n$n.json
  else
    timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --gpus $n --steps $steps --warmup 3 --workload $w 2>/dev/null | grep '^{' | tail -1 > gpurun_out/scale_${w}_n$n.json
  fi
  echo "$w n=$n: $(python tools/brief.py < gpurun_out/scale_${w}_n$n.json)"
done
[ "$2" = extra ] || exit 0
timeout 120 python - <<PY
import sys, time; sys.path.insert(0, ".")
import numpy as np
import sequenza_b200 as S
from sequenza_b200.synth import synth_segments, baf_args
seg = synth_segments(100000)
ce, pl = S.r_seq(0.1, 1, 0.001), S.r_seq(1, 7, 0.01)
for g in (1, $n):
    kw = dict(cellularity=ce, ploidy=pl, n_gpus=g, priors_table={"CN": [2], "value": [2]}, **baf_args(seg))
    S.baf_grid_sums(**kw)
    t = time.perf_counter(); r = S.baf_grid_sums(**kw); dt = time.perf_counter() - t
    print("single process, sqz_baf_grid n_gpus=%d (config 3, host vectors in/out): %.1f ms" % (g, dt * 1e3))
PY
timeout 120 python tools/bench_extra.py config5 2>/dev/null | cut -c1-400
