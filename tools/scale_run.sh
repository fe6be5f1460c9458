#!/bin/bash
# Scaling runs on one multi-GPU box: torchrun one rank per GPU (bench.py contract) for
# config 2 (weak) and config 3 (strong), then the single-process n_gpus host path.
NG=${1:-8}
port=29600
for n in 1 2 4 8; do
  [ $n -gt $NG ] && break
  for w in config2 config3; do
    steps=30; [ $w = config3 ] && steps=5
    port=$((port+1))
    if [ $n -eq 1 ]; then
      python bench.py --gpus 1 --steps $steps --warmup 3 --workload $w 2>/dev/null | tail -1 > gpurun_out/scale_${w}_n$n.json
    else
      python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --gpus $n --steps $steps --warmup 3 --workload $w 2>/dev/null | grep '^{' | tail -1 > gpurun_out/scale_${w}_n$n.json
    fi
    echo "$w n=$n: $(python tools/brief.py < gpurun_out/scale_${w}_n$n.json)"
  done
done
python - <<PY
import sys, time; sys.path.insert(0, ".")
import numpy as np
import sequenza_b200 as S
from sequenza_b200.synth import synth_segments, baf_args
seg = synth_segments(100000)
ce, pl = S.r_seq(0.1, 1, 0.001), S.r_seq(1, 7, 0.01)
ref = None
for g in (1, 2, 4, 8):
    if g > S.lib().sqz_device_count(): break
    kw = dict(cellularity=ce, ploidy=pl, n_gpus=g, priors_table={"CN": [2], "value": [2]}, **baf_args(seg))
    S.baf_grid_sums(**kw)
    t = time.perf_counter(); r = S.baf_grid_sums(**kw); dt = time.perf_counter() - t
    if ref is None: ref = r
    print("single process, sqz_baf_grid n_gpus=%d (config 3, host vectors in/out): %.1f ms, identical to n_gpus=1: %s" % (g, dt * 1e3, bool(np.array_equal(r, ref))))
PY
python tools/bench_extra.py config5 2>/dev/null | cut -c1-400
