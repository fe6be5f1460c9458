#!/usr/bin/env python3
"""Latency of the small, interactive-size calls: sequenza.fit on an example-sized segment
table (BASELINE config 1 shape), mufreq.model.fit on a few thousand mutations, baf.bayes at
the solution."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sequenza_b200 as S
from sequenza_b200.synth import baf_args, synth_mutations, synth_segments
PRI = {"CN": [2], "value": [2]}
def best(fn, rep=20):
    fn(); ts = []
    for _ in range(rep):
        t = time.perf_counter(); fn(); ts.append(time.perf_counter() - t)
    return min(ts), float(np.median(ts))
seg = synth_segments(150, seed=3)
ce, pl = S.r_seq(0.1, 1, 0.01), S.r_seq(1, 7, 0.1)
a = best(lambda: S.baf_model_fit(cellularity=ce, ploidy=pl, priors_table=PRI, **baf_args(seg)))
b = best(lambda: S.baf_grid_sums(cellularity=ce, ploidy=pl, priors_table=PRI, **baf_args(seg)))
print(json.dumps({"call": "baf.model.fit, 150 segments (example-sized), grid 91x61, CNt.max=7",
                  "ms_min": a[0]*1e3, "ms_median": a[1]*1e3, "ms_C_call_only_min": b[0]*1e3}))
c = best(lambda: S.baf_bayes(cellularity=0.7, ploidy=2.4, CNt_max=20, **baf_args(seg)))
print(json.dumps({"call": "baf.bayes, 150 segments, CNt.max=20", "ms_min": c[0]*1e3}))
mut = synth_mutations(3000, seed=4)
w = np.full(3000, 60.0)
d = best(lambda: S.mufreq_grid_sums(mut["mufreq"], mut["depth_ratio"], ce, pl, 1.0, w, 2*w, priors_table=PRI), rep=5)
print(json.dumps({"call": "mufreq.model.fit sums, 3000 mutations, grid 91x61, T=35", "ms_min": d[0]*1e3,
                  "position_x_gridpoint_per_s": 3000*ce.size*pl.size/d[0]}))
