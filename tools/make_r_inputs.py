#!/usr/bin/env python3
"""Writes the synthetic inputs that tools/r_reference_dump.R feeds to the unmodified R
package: tests/golden/r_inputs/{baf_segments.tsv, mufreq_positions.tsv, grid.json}
(%.17g, so that R reads the very doubles the tests use).  Deterministic; committed."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from sequenza_b200.bayes import r_seq  # noqa: E402
from sequenza_b200.synth import synth_mutations, synth_segments  # noqa: E402

out = os.path.join(ROOT, "tests", "golden", "r_inputs")
os.makedirs(out, exist_ok=True)
seg = synth_segments(400, seed=20190509)
keys = ("Bf", "depth_ratio", "sd_Bf", "sd_ratio", "weight_Bf", "weight_ratio")
with open(os.path.join(out, "baf_segments.tsv"), "w") as f:
    f.write("\t".join(keys) + "\n")
    for i in range(400):
        f.write("\t".join("NA" if np.isnan(seg[k][i]) else "%.17g" % seg[k][i] for k in keys) + "\n")
mut = synth_mutations(600, seed=20190509)
w = np.round(np.random.default_rng(20190509).uniform(20, 200, 600))
with open(os.path.join(out, "mufreq_positions.tsv"), "w") as f:
    f.write("mufreq\tdepth_ratio\tweight\n")
    for i in range(600):
        f.write("%.17g\t%.17g\t%.17g\n" % (mut["mufreq"][i], mut["depth_ratio"][i], w[i]))
with open(os.path.join(out, "grid.json"), "w") as f:
    json.dump({"cellularity": [float("%.17g" % v) for v in r_seq(0.1, 1, 0.05)],
               "ploidy": [float("%.17g" % v) for v in r_seq(1, 7, 0.25)],
               "point": [0.7, 2.4]}, f)
print("wrote", out)
