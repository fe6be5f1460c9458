"""Config 3 shards (the columns one of N GPUs gets): column-per-lane kernel with 16- vs
32-wide ploidy tiles, timed on one GPU.  Source of the tile-width rule in colwarp_tile."""
import os, sys
sys.path.insert(0, ".")
import torch
import sequenza_b200 as S
from sequenza_b200.dist import BafGridDevice
from sequenza_b200.synth import synth_segments, baf_args
PRI = {"CN": [2], "value": [2]}
seg = synth_segments(100000)
ce, pl = S.r_seq(0.1, 1, 0.001), S.r_seq(1, 7, 0.01)
dev = BafGridDevice(**baf_args(seg))
for ncol in (57, 113, 226, 451):
    res = []
    for pt in ("16", "32"):
        os.environ["SQZ_COLWARP"] = "1"; os.environ["SQZ_PT"] = pt
        dev.run(pl, ce, 0, ncol, priors_table=PRI); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3): dev.run(pl, ce, 0, ncol, priors_table=PRI)
        e1.record(); torch.cuda.synchronize()
        res.append(e0.elapsed_time(e1) / 3)
    groups = (ncol + 31) // 32
    print("columns %3d (%2d groups; 32-wide tiles: %5.1f waves of one-warp CTAs): PT16 %.2f ms  PT32 %.2f ms"
          % (ncol, groups, groups * 19 * 64 / 1184.0, res[0], res[1]), flush=True)

for k in ("SQZ_COLWARP", "SQZ_PT"):
    os.environ.pop(k, None)
for ncol in (113, 901):
    dev.run(pl, ce, 0, ncol, priors_table=PRI); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3): dev.run(pl, ce, 0, ncol, priors_table=PRI)
    e1.record(); torch.cuda.synchronize()
    print("columns %3d, planner's own choice: %.2f ms" % (ncol, e0.elapsed_time(e1) / 3), flush=True)
