"""Pruned vs split kernel over sd scales and grid resolutions (CNt 0..7)."""
import os, sys
sys.path.insert(0, ".")
import torch
import sequenza_b200 as S
from sequenza_b200.dist import BafGridDevice
from sequenza_b200.synth import synth_segments, baf_args
PRI = {"CN": [2], "value": [2]}
grids = {"fine 181x601": (S.r_seq(0.1, 1, 0.005), S.r_seq(1, 7, 0.01)),
         "default 91x61": (S.r_seq(0.1, 1, 0.01), S.r_seq(1, 7, 0.1)),
         "mid 91x301": (S.r_seq(0.1, 1, 0.01), S.r_seq(1, 7, 0.02))}
for gname, (ce, pl) in grids.items():
    for scale in (0.03, 0.3, 1.0, 10.0):
        seg = synth_segments(100000 if "fine" in gname else 30000, seed=3)
        seg["sd_ratio"] = seg["sd_ratio"] * scale
        seg["sd_Bf"] = seg["sd_Bf"] * scale
        dev = BafGridDevice(**baf_args(seg))
        res = {}
        for prune in ("0", "1"):
            os.environ["SQZ_PRUNE"] = prune
            for pt in ("", "16", "8"):
                if pt: os.environ["SQZ_PT"] = pt
                else: os.environ.pop("SQZ_PT", None)
                out = dev.run(pl, ce, priors_table=PRI); torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(3): out = dev.run(pl, ce, priors_table=PRI)
                e1.record(); torch.cuda.synchronize()
                res[(prune, pt)] = e0.elapsed_time(e1) / 3
        os.environ.pop("SQZ_PT", None)
        print("%-14s sd x%-5g split %.3f | pruned PT32 %.3f PT16 %.3f PT8 %.3f ms" % (
            gname, scale, res[("0", "")], res[("1", "")], res[("1", "16")], res[("1", "8")]))
