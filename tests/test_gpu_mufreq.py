"""Parity of the CUDA mufreq path (sqz_mufreq_point, sqz_mufreq_grid) with the oracle."""
import numpy as np
import pytest

import sequenza_b200 as S
from sequenza_b200.synth import synth_mutations

pytestmark = pytest.mark.gpu
RTOL = 1e-9
PRI = {"CN": [2], "value": [2]}


def same(a, b):
    return bool(np.all((a == b) | (np.isnan(a) & np.isnan(b))))


@pytest.mark.parametrize("cn_max", [7, 20])
def test_mufreq_bayes_scalar_weights(oracle, cn_max):
    """As sequenza.results calls it (R/results.R:153-157): default weights 100 / 100."""
    mut = synth_mutations(20000, seed=3 + cn_max)
    for cell, plo in ((0.7, 2.4), (1.0, 2.0), (0.15, 5.5)):
        a = S.mufreq_bayes(mut["mufreq"], mut["depth_ratio"], cell, plo, 1.0, CNt_max=cn_max)
        b = oracle.mufreq_bayes(mut["mufreq"], mut["depth_ratio"], cell, plo, 1.0,
                                CNt_max=cn_max)
        for col in range(3):
            assert same(a[:, col], b[:, col]), (cell, plo, col)
        np.testing.assert_allclose(a[:, 3], b[:, 3], rtol=RTOL)


def test_mufreq_bayes_vector_weights_and_priors(oracle):
    """As sequenza.fit(method = "mufreq") weights it (R/fit.R:60-65)."""
    mut = synth_mutations(5000, seed=8)
    rng = np.random.default_rng(2)
    w = rng.integers(8, 700, 5000).astype(float)
    f = np.rint(mut["mufreq"] * w) / w
    a = S.mufreq_bayes(f, mut["depth_ratio"], 0.55, 3.2, 1.02, w, 2 * w, priors_table=PRI)
    b = oracle.mufreq_bayes(f, mut["depth_ratio"], 0.55, 3.2, 1.02, w, 2 * w,
                            priors_table=PRI)
    for col in range(3):
        assert same(a[:, col], b[:, col])
    np.testing.assert_allclose(a[:, 3], b[:, 3], rtol=RTOL)


def test_mufreq_bayes_edges(oracle):
    """F = 0 and 1, underflow to the 1e-323 floor, NA inputs, non-integer depth."""
    f = np.array([0.0, 1.0, 0.5, 0.999, np.nan, 0.3, 0.3])
    r = np.array([1.0, 1.0, 50.0, 1e-3, 1.0, np.nan, 1.0])
    wm = np.array([100.0, 100, 5000, 5000, 100, 100, 10.5])
    a = S.mufreq_bayes(f, r, 0.9, 2.0, 1.0, wm, 2 * wm)
    b = oracle.mufreq_bayes(f, r, 0.9, 2.0, 1.0, wm, 2 * wm)
    for col in range(4):
        g, h = a[:, col], b[:, col]
        ok = (g == h) | (np.isnan(g) & np.isnan(h)) | (np.abs(g - h) <= RTOL * np.abs(h))
        assert ok.all(), (col, a, b)
    assert S.mufreq_bayes([], [], 0.5, 2.0, 1.0).shape == (0, 4)


def test_mufreq_grid_matches_oracle(oracle):
    mut = synth_mutations(800, seed=21)
    w = np.random.default_rng(4).integers(20, 200, 800).astype(float)
    ce, pl = S.r_seq(0.2, 1, 0.05), S.r_seq(1, 5, 0.25)
    kw = dict(cellularity=ce, ploidy=pl, avg_depth_ratio=1.0, weight_mufreq=w,
              weight_ratio=2 * w, priors_table=PRI)
    got = S.mufreq_grid_sums(mut["mufreq"], mut["depth_ratio"], **kw)
    ref = oracle.mufreq_grid_sums(mut["mufreq"], mut["depth_ratio"], **kw)
    np.testing.assert_allclose(got, ref, rtol=RTOL)
    assert int(np.argmax(got)) == int(np.argmax(ref))
    a = S.mufreq_model_fit(mufreq=mut["mufreq"], depth_ratio=mut["depth_ratio"], **kw)
    b = oracle.finish_model_fit(ref, pl, ce)
    big = b["lpp"] > 1e-290
    np.testing.assert_allclose(a["lpp"][big], b["lpp"][big], rtol=RTOL)


def test_mufreq_scalar_table_path_equals_vector_path(oracle):
    """Scalar weights take the table-driven kernel (sqz_mufreq_point_scalar); the same call
    with the weights spelled out as vectors takes the direct kernel.  Both must give the
    oracle's rows, including positions that fall outside the table (NA, ratio < 0, F > 1)."""
    mut = synth_mutations(30000, seed=17)
    f, r = mut["mufreq"].copy(), mut["depth_ratio"].copy()
    f[:6] = [np.nan, 0.0, 1.0, 1.2, -0.1, 0.5]
    r[:6] = [1.0, np.nan, -0.5, 1.0, 1.0, 1e300]
    for wm, wr, cn_max in ((100, 100, 7), (100, 100, 20), (37, 250, 7), (1, 1, 7)):
        a = S.mufreq_bayes(f, r, 0.7, 2.4, 1.0, wm, wr, CNt_max=cn_max, priors_table=PRI)
        b = S.mufreq_bayes(f, r, 0.7, 2.4, 1.0, np.full(f.size, float(wm)),
                           np.full(f.size, float(wr)), CNt_max=cn_max, priors_table=PRI)
        c = oracle.mufreq_bayes(f, r, 0.7, 2.4, 1.0, wm, wr, CNt_max=cn_max, priors_table=PRI)
        for col in range(4):
            assert same(a[:, col], b[:, col]), (wm, wr, col)
        for col in range(3):
            assert same(a[:, col], c[:, col]), (wm, wr, col)
        ok = (a[:, 3] == c[:, 3]) | (np.isnan(a[:, 3]) & np.isnan(c[:, 3])) | \
             (np.abs(a[:, 3] - c[:, 3]) <= RTOL * np.abs(c[:, 3]))
        assert ok.all()
    # non-integer scalar weight: no table, direct path, dbinom gives NaN as in R
    a = S.mufreq_bayes(f[6:50], r[6:50], 0.7, 2.4, 1.0, 10.5, 100)
    c = oracle.mufreq_bayes(f[6:50], r[6:50], 0.7, 2.4, 1.0, 10.5, 100)
    assert same(a[:, 1], c[:, 1]) and same(a[:, 3], c[:, 3])


@pytest.mark.parametrize("n", [70000, 70001])
def test_mufreq_scalar_table_in_shared_memory(oracle, n):
    """From 65,536 positions on (even count, aligned vectors) the lookup keeps the compact
    answer table in shared memory (mufreq_lookup_smem_kernel); an odd count takes the
    global-memory lookup.  Both against the direct kernel on every row -- positions outside
    the table spread over the vector so that every thread role meets some -- and against the
    oracle on a sample."""
    mut = synth_mutations(n, seed=23)
    f, r = mut["mufreq"].copy(), mut["depth_ratio"].copy()
    rng = np.random.default_rng(5)
    idx = rng.choice(n, 400, replace=False)
    f[idx[:100]] = np.nan
    r[idx[100:200]] = np.nan
    f[idx[200:300]] = rng.choice([1.2, -0.1, 7.0], 100)
    r[idx[300:400]] = rng.choice([-0.5, -2.0, 1e300], 100)
    for wm, wr, cn_max in ((100, 100, 20), (100, 100, 7), (37, 250, 7)):
        a = S.mufreq_bayes(f, r, 0.7, 2.4, 1.0, wm, wr, CNt_max=cn_max, priors_table=PRI)
        b = S.mufreq_bayes(f, r, 0.7, 2.4, 1.0, np.full(n, float(wm)), np.full(n, float(wr)),
                           CNt_max=cn_max, priors_table=PRI)
        for col in range(4):
            assert same(a[:, col], b[:, col]), (wm, wr, col)
        sub = np.concatenate([idx, np.arange(0, n, 97)])
        c = oracle.mufreq_bayes(f[sub], r[sub], 0.7, 2.4, 1.0, wm, wr, CNt_max=cn_max,
                                priors_table=PRI)
        for col in range(3):
            assert same(a[sub, col], c[:, col]), (wm, wr, col)
        ok = (a[sub, 3] == c[:, 3]) | (np.isnan(a[sub, 3]) & np.isnan(c[:, 3])) | \
             (np.abs(a[sub, 3] - c[:, 3]) <= RTOL * np.abs(c[:, 3]))
        assert ok.all()
