"""sqz_fit_finish: normalisation (R/bayes.R:120-122) and the get.ci reductions
(R/solutions.R:1-15) on the device, against the step-by-step host versions and the oracle."""
import numpy as np
import pytest

import sequenza_b200 as S
from sequenza_b200.synth import baf_args, synth_segments

pytestmark = pytest.mark.gpu
PRI = {"CN": [2], "value": [2]}


def check_fit(sums, pl, ce, oracle, levels=(0.95, 0.5, 0.999, 0.05)):
    dev = S.finish_model_fit(sums, pl, ce, device=True)
    host = S.finish_model_fit(sums, pl, ce, device=False)
    ref = oracle.finish_model_fit(sums, pl, ce)
    assert dev.ci_cache is not None and getattr(host, "ci_cache", None) is None
    np.testing.assert_array_equal(dev["ploidy"], host["ploidy"])
    np.testing.assert_array_equal(dev["cellularity"], host["cellularity"])
    assert dev["lpp"].shape == host["lpp"].shape == (len(pl), len(ce))
    big = ref["lpp"] > 1e-290
    np.testing.assert_allclose(dev["lpp"][big], ref["lpp"][big], rtol=1e-9, atol=0)
    np.testing.assert_allclose(dev["lpp"], host["lpp"], rtol=1e-12, atol=1e-300)
    assert abs(dev["lpp"].sum() - 1) < 1e-9
    for lvl in levels:
        a, b, c = S.get_ci(dev, lvl), S.get_ci(host, lvl), oracle.get_ci(ref, lvl)
        for key in ("max.ploidy", "max.cellularity"):
            assert a[key] == b[key] == c[key]
        for key in ("confint.ploidy", "confint.cellularity"):
            assert a[key] == b[key] == tuple(c[key]), (lvl, key)
        np.testing.assert_allclose(a["values.ploidy"]["y"], c["values.ploidy"]["y"], rtol=1e-9)
        np.testing.assert_allclose(a["values.cellularity"]["x"], c["values.cellularity"]["x"],
                                   rtol=1e-9)
    return dev


def test_finish_matches_host_and_oracle_default_grid(oracle):
    ce, pl = S.r_seq(0.1, 1, 0.01), S.r_seq(1, 7, 0.1)
    for seed, n in ((1, 40), (2, 400), (3, 3000)):  # broad, medium and sharp posteriors
        seg = synth_segments(n, seed=seed)
        sums = S.baf_grid_sums(cellularity=ce, ploidy=pl, priors_table=PRI, **baf_args(seg))
        check_fit(sums, pl, ce, oracle)


def test_finish_ties_flat_and_tiny(oracle):
    rng = np.random.default_rng(0)
    pl, ce = np.array([1.0, 2.0, 3.0, 4.0]), np.array([0.2, 0.5, 0.9])
    check_fit(np.full(12, -100.0), pl, ce, oracle)                       # flat: all tied
    s = -50.0 - rng.integers(0, 3, 12).astype(float)                      # many exact ties
    check_fit(s, pl, ce, oracle)
    check_fit(np.array([-3.0]), np.array([2.0]), np.array([0.5]), oracle)  # 1 x 1 grid
    big = -7e5 + 200 * rng.standard_normal(37 * 53)                       # underflowing tails
    check_fit(big, np.arange(37) * 0.1 + 1, np.arange(53) * 0.01 + 0.1, oracle)


def test_finish_fallback_to_host_steps():
    # unsorted / duplicated axes and NaN keep the reference's host steps (tapply semantics)
    pl, ce = np.array([2.0, 1.0, 2.0]), np.array([0.5, 0.25])
    sums = np.array([-10.0, -11.0, -12.0, -9.0, -13.0, -9.5])
    t = S.finish_model_fit(sums, pl, ce)
    assert getattr(t, "ci_cache", None) is None and t["lpp"].shape == (2, 2)
    with pytest.raises(ValueError):
        S.finish_model_fit(sums, pl, ce, device=True)
    s2 = np.array([-1.0, np.nan, -2.0, -3.0])
    t = S.finish_model_fit(s2, [1.0, 2.0], [0.3, 0.6])
    assert getattr(t, "ci_cache", None) is None and np.isnan(t["lpp"]).all()


def test_finish_fine_grid_and_cohort(oracle):
    """901 x 601 (BASELINE config 3's grid) and a 16-sample cohort in one launch."""
    ce, pl = S.r_seq(0.1, 1, 0.001), S.r_seq(1, 7, 0.01)
    seg = synth_segments(2000, seed=5)
    sums = S.baf_grid_sums(cellularity=ce, ploidy=pl, priors_table=PRI, **baf_args(seg))
    dev = S.finish_model_fit(sums, pl, ce, device=True)
    host = S.finish_model_fit(sums, pl, ce, device=False)
    a, b = S.get_ci(dev), S.get_ci(host)
    assert a["confint.ploidy"] == b["confint.ploidy"]
    assert a["confint.cellularity"] == b["confint.cellularity"]
    assert a["max.ploidy"] == b["max.ploidy"] and a["max.cellularity"] == b["max.cellularity"]
    np.testing.assert_allclose(dev["lpp"], host["lpp"], rtol=1e-12, atol=1e-300)
    ce, pl = S.r_seq(0.1, 1, 0.02), S.r_seq(1, 7, 0.2)
    samples = [baf_args(synth_segments(100 + 20 * k, cellularity=0.3 + 0.04 * k, seed=70 + k))
               for k in range(16)]
    fits = S.baf_model_fit_batch(samples, cellularity=ce, ploidy=pl, priors_table=PRI)
    for smp, f in zip(samples, fits):
        one = S.finish_model_fit(S.baf_grid_sums(cellularity=ce, ploidy=pl, priors_table=PRI,
                                                 **smp), pl, ce, device=False)
        np.testing.assert_allclose(f["lpp"], one["lpp"], rtol=1e-12, atol=1e-300)
        assert S.get_ci(f)["confint.cellularity"] == S.get_ci(one)["confint.cellularity"]
        assert S.get_ci(f)["confint.ploidy"] == S.get_ci(one)["confint.ploidy"]


def test_finish_small_grid_many_samples_repeated():
    """A 5 x 7 grid for 512 samples, 20 times: on such a small grid most threads of a CTA
    have no grid point and race ahead -- the shape in which a missing barrier between the
    maximum and the sum reductions would show as an occasional wrong LogSumLik."""
    rng = np.random.default_rng(3)
    pl, ce = np.arange(5) * 0.5 + 1.0, np.arange(7) * 0.1 + 0.2
    sums = -500.0 + 30.0 * rng.standard_normal((512, 35))
    want = [S.finish_model_fit(sums[k], pl, ce, device=False)["lpp"] for k in range(512)]
    from sequenza_b200.bayes import finish_model_fit_device
    for _ in range(20):
        got = finish_model_fit_device(sums, pl, ce)
        for k in range(512):
            np.testing.assert_allclose(got[k]["lpp"], want[k], rtol=1e-12, atol=1e-300)
