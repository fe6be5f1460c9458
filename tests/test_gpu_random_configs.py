"""Differential test over randomly drawn call configurations (fixed seeds): copy-number
ranges that do not start at 0, CNn = 1 / 2 / 3, arbitrary priors tables (including zero
priors and repeated CN rows), extreme sd and weights, NA patterns, both selection rules,
odd grid shapes.  GPU entry points against the oracle."""
import numpy as np
import pytest

import sequenza_b200 as S

pytestmark = pytest.mark.gpu
RTOL = 1e-9


def rel_close(a, b, rtol=RTOL):
    a, b = np.asarray(a), np.asarray(b)
    with np.errstate(invalid="ignore"):
        ok = (np.abs(a - b) <= rtol * np.abs(b)) | (np.isnan(a) & np.isnan(b)) | \
             (np.isinf(a) & (a == b))
    return bool(ok.all())


def draw(rng):
    n = int(rng.integers(1, 400))
    cn_min = int(rng.integers(0, 3))
    cn_max = cn_min + int(rng.integers(0, 12))
    cnn = int(rng.choice([2, 2, 2, 1, 3]))
    Bf = rng.uniform(0, 0.5, n)
    Bf[rng.random(n) < rng.choice([0.0, 0.05, 0.5])] = np.nan
    if cnn < 2:
        Bf[:] = np.nan  # the reference only calls CNn = 1 with Bf = NA (R/results.R:139-145)
    ratio = np.exp(rng.normal(0, rng.choice([0.1, 0.5, 1.5]), n))
    sd_scale = rng.choice([1e-3, 0.1, 1.0, 30.0])
    kw = dict(
        Bf=Bf, depth_ratio=ratio,
        sd_Bf=rng.uniform(0.01, 0.2, n) * rng.choice([1.0, 10.0]),
        sd_ratio=rng.uniform(0.2, 1.5, n) * sd_scale,
        weight_Bf=rng.choice([1.0, 7.5]) if rng.random() < 0.3 else rng.uniform(1, 200, n),
        weight_ratio=rng.uniform(1, 200, n),
        avg_depth_ratio=float(rng.choice([1.0, 0.93, 1.2])),
        CNt_min=cn_min, CNt_max=cn_max, CNn=cnn,
        ratio_priority=bool(rng.integers(0, 2)),
    )
    k = int(rng.integers(0, 4))
    if k:
        cn = rng.integers(cn_min, cn_max + 1, k).astype(float)
        val = rng.choice([0.0, 0.5, 1.0, 2.0, 10.0], k)
        if np.all(val == 0) and cn_max == cn_min:
            val[0] = 1.0
        kw["priors_table"] = {"CN": cn.tolist(), "value": val.tolist()}
    return kw


@pytest.mark.parametrize("seed", range(24))
def test_random_configuration(oracle, seed):
    rng = np.random.default_rng(1000 + seed)
    kw = draw(rng)
    cell = float(rng.choice([0.05, 0.3, 0.77, 1.0]))
    plo = float(rng.uniform(0.8, 7.5))
    a = S.baf_bayes(cellularity=cell, ploidy=plo, **kw)
    b, margin = oracle.baf_bayes(cellularity=cell, ploidy=plo, return_margin=True, **kw)
    tie = (margin > 0) & (margin < 1e-12)
    for col in range(3):
        same = (a[:, col] == b[:, col]) | (np.isnan(a[:, col]) & np.isnan(b[:, col]))
        assert np.all(same | tie), (seed, col)
    assert rel_close(a[:, 3], b[:, 3]), seed
    nC, nP = int(rng.integers(1, 9)), int(rng.integers(1, 40))
    ce = np.sort(rng.uniform(0.02, 1.0, nC))
    if rng.random() < 0.3:
        ce[-1] = 1.0
    pl = np.sort(rng.uniform(0.5, 8.0, nP))
    got = S.baf_grid_sums(cellularity=ce, ploidy=pl, **kw)
    ref = oracle.baf_grid_sums(cellularity=ce, ploidy=pl, **kw)
    assert rel_close(got, ref), seed
    fin = np.isfinite(ref)
    if fin.any() and np.sum(ref[fin] == ref[fin].max()) == 1:
        assert int(np.nanargmax(np.where(fin, got, -np.inf))) == \
            int(np.nanargmax(np.where(fin, ref, -np.inf)))


@pytest.mark.parametrize("seed", range(8))
def test_random_mufreq_configuration(oracle, seed):
    rng = np.random.default_rng(2000 + seed)
    n = int(rng.integers(1, 300))
    cn_min = int(rng.integers(1, 3))
    cn_max = cn_min + int(rng.integers(0, 10))
    depth = rng.integers(5, 400, n).astype(float)
    f = np.rint(rng.random(n) * depth) / depth
    r = np.exp(rng.normal(0, 0.6, n))
    scalar = bool(rng.integers(0, 2))
    wm = float(rng.integers(5, 300)) if scalar else depth
    wr = float(rng.integers(5, 300)) if scalar else 2 * depth
    kw = dict(CNt_min=cn_min, CNt_max=cn_max, CNn=int(rng.choice([2, 2, 1])))
    if rng.random() < 0.5:
        kw["priors_table"] = {"CN": [float(cn_min), 2.0], "value": [3.0, 2.0]}
    cell, plo = float(rng.choice([0.2, 0.66, 1.0])), float(rng.uniform(1, 6))
    a = S.mufreq_bayes(f, r, cell, plo, 1.0, wm, wr, **kw)
    b = oracle.mufreq_bayes(f, r, cell, plo, 1.0, wm, wr, **kw)
    for col in range(3):
        assert bool(np.all((a[:, col] == b[:, col]) | (np.isnan(a[:, col]) & np.isnan(b[:, col]))))
    assert rel_close(a[:, 3], b[:, 3])
    ce, pl = np.sort(rng.uniform(0.1, 1, 3)), np.sort(rng.uniform(1, 6, 4))
    got = S.mufreq_grid_sums(f, r, ce, pl, 1.0, wm, wr, **kw)
    ref = oracle.mufreq_grid_sums(f, r, ce, pl, 1.0, wm, wr, **kw)
    assert rel_close(got, ref)
