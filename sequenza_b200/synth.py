"""Synthetic seqz-derived inputs of the shapes BASELINE.json names: a
segment table for baf.model.fit and an SNV table for mufreq.bayes, drawn around a known
(cellularity, ploidy).  numpy Generator(PCG64(seed)); default seed 20190509."""
from __future__ import annotations

import numpy as np

from .model import theoretical_baf, theoretical_depth_ratio, theoretical_mufreq

DEFAULT_SEED = 20190509


def _copy_numbers(rng, n, ploidy, cn_max):
    cn = np.arange(0, cn_max + 1)
    pmf = np.exp(-0.9 * np.abs(cn - round(ploidy)))
    pmf[0] *= 0.25  # homozygous deletions are rare
    return rng.choice(cn, size=n, p=pmf / pmf.sum())


def synth_segments(n, cellularity=0.7, ploidy=2.4, cn_max=7, seed=DEFAULT_SEED,
                   na_frac=0.01, avg_depth_ratio=1.0):
    """Segment table as sequenza.fit hands it to baf.model.fit (R/fit.R:45-50):
    weights are segment lengths in Mb for both terms."""
    rng = np.random.Generator(np.random.PCG64(seed))
    cnt = _copy_numbers(rng, n, ploidy, cn_max)
    b = np.floor(rng.random(n) * (cnt // 2 + 1))
    ratio = theoretical_depth_ratio(cnt, cellularity, ploidy,
                                    avg_depth_ratio=avg_depth_ratio)
    ratio = np.maximum(ratio * np.exp(rng.normal(0.0, 0.05, n)), 1e-3)
    bf = np.clip(theoretical_baf(cnt, b, cellularity) + rng.normal(0.0, 0.02, n), 0.0, 0.5)
    if na_frac > 0:
        bf[rng.random(n) < na_frac] = np.nan
    length = np.exp(rng.uniform(np.log(3e6), np.log(1e8), n))
    return {
        "Bf": bf,
        "depth_ratio": ratio,
        "sd_Bf": rng.uniform(0.03, 0.12, n),
        "sd_ratio": rng.uniform(0.3, 1.5, n),
        "weight_Bf": length / 1e6,
        "weight_ratio": length / 1e6,
        "avg_depth_ratio": avg_depth_ratio,
        "truth": {"cellularity": cellularity, "ploidy": ploidy, "CNt": cnt, "B": b},
    }


def synth_mutations(n, cellularity=0.7, ploidy=2.4, cn_max=7, seed=DEFAULT_SEED, depth=100,
                    avg_depth_ratio=1.0):
    """Per-position SNV table for mufreq.bayes: F from binomial noise at `depth` around
    theoretical.mufreq, depth ratio of the containing segment with lognormal noise."""
    rng = np.random.Generator(np.random.PCG64(seed + 1))
    cnt = np.maximum(_copy_numbers(rng, n, ploidy, cn_max), 1)
    mt = np.floor(rng.random(n) * (cnt + 1))
    mf = np.clip(theoretical_mufreq(cnt, mt, cellularity), 0.01, 1.0)
    f = rng.binomial(depth, mf) / float(depth)
    ratio = theoretical_depth_ratio(cnt, cellularity, ploidy,
                                    avg_depth_ratio=avg_depth_ratio)
    ratio = np.maximum(ratio * np.exp(rng.normal(0.0, 0.08, n)), 1e-3)
    return {"mufreq": f, "depth_ratio": ratio, "avg_depth_ratio": avg_depth_ratio,
            "truth": {"cellularity": cellularity, "ploidy": ploidy, "CNt": cnt, "Mt": mt}}


BAF_KEYS = ("Bf", "depth_ratio", "sd_Bf", "sd_ratio", "weight_Bf", "weight_ratio",
            "avg_depth_ratio")


def baf_args(seg):
    """The baf.model.fit / baf.bayes keyword arguments of a synth_segments() table."""
    return {k: seg[k] for k in BAF_KEYS}
