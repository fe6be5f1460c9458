"""Column-per-lane kernel (SQZ_COLWARP=1) against the segment-per-lane pruned kernel
(SQZ_COLWARP=0) over grid shapes, segment counts, sd scales and column counts (CNt 0..7)."""
import os, sys
import numpy as np
sys.path.insert(0, ".")
import torch
import sequenza_b200 as S
from sequenza_b200.dist import BafGridDevice
from sequenza_b200.synth import synth_segments, baf_args
PRI = {"CN": [2], "value": [2]}


CN_MAX = int(sys.argv[1]) if len(sys.argv) > 1 else 7  # 20: the other kernel is the split one


def timed(dev, pl, ce, env):
    for k in ("SQZ_COLWARP", "SQZ_PT"):
        os.environ.pop(k, None)
    os.environ.update(env)
    dev.run(pl, ce, priors_table=PRI, CNt_max=CN_MAX); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3): dev.run(pl, ce, priors_table=PRI, CNt_max=CN_MAX)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / 3


cases = [("default 91x61", S.r_seq(0.1, 1, 0.01), S.r_seq(1, 7, 0.1)),
         ("mid 91x301", S.r_seq(0.1, 1, 0.01), S.r_seq(1, 7, 0.02)),
         ("fine 181x601", S.r_seq(0.1, 1, 0.005), S.r_seq(1, 7, 0.01))]
for name, ce, pl in cases:
    for n in (300, 3000, 30000):
        for scale in (0.1, 1.0, 10.0):
            if scale != 1.0 and n != 30000:
                continue
            seg = synth_segments(n, seed=3, cn_max=CN_MAX)
            seg["sd_ratio"] = seg["sd_ratio"] * scale
            seg["sd_Bf"] = seg["sd_Bf"] * scale
            dev = BafGridDevice(**baf_args(seg))
            print("%-14s n=%-6d sd x%-4g pruned %.3f | column-lane PT16 %.3f PT32 %.3f ms" % (
                name, n, scale, timed(dev, pl, ce, {"SQZ_COLWARP": "0"}),
                timed(dev, pl, ce, {"SQZ_COLWARP": "1", "SQZ_PT": "16"}),
                timed(dev, pl, ce, {"SQZ_COLWARP": "1", "SQZ_PT": "32"})), flush=True)
# lane efficiency: column counts that leave part of the last 32-column group idle
seg = synth_segments(30000, seed=3, cn_max=CN_MAX)
dev = BafGridDevice(**baf_args(seg))
pl = S.r_seq(1, 7, 0.02)
for nc in (8, 12, 16, 20, 24, 33, 40, 48):
    ce = np.linspace(0.1, 1.0, nc)
    print("columns %-3d (301 ploidies, n=30000): pruned %.3f | column-lane %.3f ms" % (
        nc, timed(dev, pl, ce, {"SQZ_COLWARP": "0"}), timed(dev, pl, ce, {"SQZ_COLWARP": "1"})),
        flush=True)
