#!/usr/bin/env python3
"""Touches every kernel once on small inputs (for compute-sanitizer runs)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sequenza_b200 as S
from sequenza_b200.synth import baf_args, synth_mutations, synth_segments
PRI = {"CN": [2], "value": [2]}
seg = synth_segments(700, seed=1)
seg["sd_ratio"][5] = 0.0  # forces the repair kernel
ce, pl = S.r_seq(0.1, 1, 0.1), S.r_seq(1, 7, 0.13)
for env in ({}, {"SQZ_SPLIT": "0"}, {"SQZ_PRUNE": "0"}, {"SQZ_PRUNE": "1", "SQZ_PT": "16"},
            {"SQZ_PRUNE": "1", "SQZ_PT": "8"}):
    for k in ("SQZ_SPLIT", "SQZ_PRUNE", "SQZ_PT"):
        os.environ.pop(k, None)
    os.environ.update(env)
    for cn, rp in ((7, False), (20, False), (7, True), (20, True)):
        r = S.baf_grid_sums(cellularity=ce, ploidy=pl, CNt_max=cn, ratio_priority=rp,
                            priors_table=PRI, **baf_args(seg))
        assert r.shape == (ce.size * pl.size,)
for k in ("SQZ_SPLIT", "SQZ_PRUNE", "SQZ_PT"):
    os.environ.pop(k, None)
S.baf_grid_sums(cellularity=ce, ploidy=pl, mode=S.SQZ_MODE_FAITHFUL, **baf_args(seg))
S.baf_model_fit_batch([baf_args(synth_segments(n, seed=n)) for n in (5, 130, 1)],
                      cellularity=ce, ploidy=pl)
S.baf_bayes(cellularity=0.5, ploidy=2.0, CNt_max=20, **baf_args(seg))
mut = synth_mutations(900)
S.mufreq_bayes(mut["mufreq"], mut["depth_ratio"], 0.5, 2.0, 1.0, CNt_max=20)
S.mufreq_bayes(mut["mufreq"], mut["depth_ratio"], 0.5, 2.0, 1.0, np.full(900, 40.0), 80.0)
S.mufreq_grid_sums(mut["mufreq"][:100], mut["depth_ratio"][:100], ce[:3], pl[:5], 1.0)
S.device_model_points("baf", 0.5, 2.0)
import torch
from sequenza_b200.dist import BafGridDevice, run_noalloc
dev = BafGridDevice(**baf_args(seg))
run_noalloc(dev, pl, ce, priors_table=PRI)
torch.cuda.synchronize()
S.lib().sqz_release()
print("all kernels ran")
