"""Pins the CPU oracle (oracle/sqz_oracle.c) against every known answer the reference
tree holds for this path, and against independent implementations (scipy, mpmath).
CPU only."""
import numpy as np
import pytest

scipy_stats = pytest.importorskip("scipy.stats")


def test_theoretical_depth_ratio_testthat(oracle, golden):
    g = golden("testthat_model.json")["theoretical_depth_ratio"]  # test.model.R:4-11
    for c in g["cases"]:
        got = oracle.theoretical_depth_ratio(c["CNt"], c["cellularity"], c["ploidy"])
        assert got == pytest.approx(c["expect"], rel=1.5e-8)  # expect_equal tolerance


def test_theoretical_baf_testthat(oracle, golden):
    g = golden("testthat_model.json")["theoretical_baf"]  # test.model.R:13-20
    for c in g["cases"]:
        got = oracle.theoretical_baf(c["CNt"], c["B"], c["cellularity"])
        if c["round"] is not None:
            got = round(got, c["round"])
        assert got == pytest.approx(c["expect"], rel=1.5e-8)


def test_theoretical_mufreq_testthat(oracle, golden):
    g = golden("testthat_model.json")["theoretical_mufreq"]  # test.model.R:22-29
    for c in g["cases"]:
        got = oracle.theoretical_mufreq(c["CNt"], c["Mt"], c["cellularity"])
        if c["round"] is not None:
            got = round(got, c["round"])
        assert got == pytest.approx(c["expect"], abs=1.5e-8)


def test_types_matrices_testthat(oracle, golden):
    g = golden("testthat_model.json")
    b = g["baf_types_matrix"]  # test.model.R:31-45
    m = oracle.baf_types_matrix(**b["args"])
    assert m[:, 0].tolist() == b["CNn"] and m[:, 1].tolist() == b["CNt"]
    assert m[:, 2].tolist() == b["B"]
    u = g["mufreq_types_matrix"]  # test.model.R:47-55
    m = oracle.mufreq_types_matrix(**u["args"])
    assert m[:, 0].tolist() == u["CNn"] and m[:, 1].tolist() == u["CNt"]
    assert m[:, 2].tolist() == u["Mt"]


def test_type_counts(oracle):
    # T = 20 (0..7), 121 (0..20) for baf; 35 (1..7), 230 (1..20) for mufreq
    assert oracle.baf_types_matrix(0, 7).shape[0] == 20
    assert oracle.baf_types_matrix(0, 20).shape[0] == 121
    assert oracle.baf_types_matrix(0, 7, 1).shape[0] == 8  # CNn < 2: B = 0 only
    assert oracle.mufreq_types_matrix(1, 7).shape[0] == 35
    assert oracle.mufreq_types_matrix(1, 20).shape[0] == 230


def vignette_table(golden):
    v = golden("vignette_rows.json")
    cols = v["columns"]
    rows = np.array([[float(x) for x in r] for r in v["rows"]])
    return {c: rows[:, i] for i, c in enumerate(cols)}


def test_vignette_rows(oracle, golden):
    """Six baf.bayes rows rendered in inst/doc/sequenza.html (3 dp).  Called as
    sequenza.results does (R/results.R:128-135) at the CP.example solution."""
    t = vignette_table(golden)
    out = oracle.baf_bayes(
        Bf=t["Bf"], depth_ratio=t["depth.ratio"], cellularity=0.88, ploidy=2.0,
        avg_depth_ratio=1.0, sd_Bf=t["sd.BAF"], sd_ratio=t["sd.ratio"], weight_Bf=1,
        weight_ratio=(t["end.pos"] - t["start.pos"]) / 1e6, CNt_max=20, CNn=2)
    assert out[:, 0].tolist() == t["CNt"].tolist()
    assert out[:, 1].tolist() == t["A"].tolist()
    assert out[:, 2].tolist() == t["B"].tolist()
    # inputs are rounded to 3 dp and avg.depth.ratio of that run is not recorded
    assert np.max(np.abs(out[:, 3] - t["LPP"])) < 5e-3


def test_cp_example_layout_and_point_estimate(oracle, golden):
    """data/CP.example.RData: rows = ploidy, columns = cellularity, normalised; the
    get.ci point estimate the vignette reports (ploidy 2, cellularity 0.88)."""
    d = golden("cp_example.json")
    assert d["names"] == ["ploidy", "cellularity", "lpp"]
    nr, nc = d["lpp_dim"]
    assert (nr, nc) == (len(d["ploidy"]), len(d["cellularity"])) == (61, 91)
    lpp = np.array(d["lpp_column_major"]).reshape(nc, nr).T
    assert abs(lpp.sum() - 1.0) < 1e-12
    assert [float(s) for s in d["lpp_rownames"]] == d["ploidy"]
    assert [float(s) for s in d["lpp_colnames"]] == d["cellularity"]
    ci = oracle.get_ci({"ploidy": d["ploidy"], "cellularity": d["cellularity"], "lpp": lpp})
    assert ci["max.ploidy"] == 2.0 and ci["max.cellularity"] == 0.88
    assert ci["confint.ploidy"][0] <= 2.0 <= ci["confint.ploidy"][1]
    assert ci["confint.cellularity"][0] <= 0.88 <= ci["confint.cellularity"][1]


def test_dt_against_scipy_and_mpmath(oracle):
    mp = pytest.importorskip("mpmath")
    mp.mp.dps = 40
    xs = np.concatenate([np.linspace(-40, 40, 321), [1e-9, 1e-4, 0.999, 1.0, 1.001, 1e4,
                                                     1e8, 3e8, 1e12, 1e150]])
    for x in xs:
        got = oracle.dt(x, 5, True)
        xm = mp.mpf(float(x))
        ref = (mp.loggamma(3) - mp.loggamma(mp.mpf(5) / 2) - mp.log(5 * mp.pi) / 2
               - 3 * mp.log1p(xm * xm / 5))
        assert abs(got - float(ref)) <= 4e-16 * max(1.0, abs(float(ref)))
        if abs(x) < 1e8:
            assert got == pytest.approx(scipy_stats.t.logpdf(x, 5), rel=1e-13, abs=1e-13)
    assert oracle.dt(np.inf, 5, True) == -np.inf
    assert np.isnan(oracle.dt(np.nan, 5, True))
    assert oracle.dt(0.3, 5, False) == pytest.approx(scipy_stats.t.pdf(0.3, 5), rel=1e-14)


def test_dbinom_against_scipy(oracle):
    worst = 0.0
    for n in (1, 2, 7, 15, 16, 30, 36, 81, 100, 137, 200, 501, 2000):
        for p in (0.0, 0.01, 0.05, 0.3, 0.5, 0.77, 0.99, 1.0):
            for k in sorted({0, 1, n // 3, n // 2, n - 1, n}):
                a, b = oracle.dbinom(k, n, p), scipy_stats.binom.pmf(k, n, p)
                if b > 1e-290:
                    worst = max(worst, abs(a - b) / b)
                else:
                    assert a < 1e-280
    assert worst < 5e-12
    assert oracle.dbinom(3, 10, 0.3, True) == pytest.approx(
        scipy_stats.binom.logpmf(3, 10, 0.3), rel=1e-13)
    assert np.isnan(oracle.dbinom(3, 10.5, 0.3))       # non-integer size -> NaN
    assert oracle.dbinom(3.4, 10, 0.3) == 0.0           # non-integer x -> 0 (R warns)
    assert oracle.dbinom(-1, 10, 0.3) == 0.0


def test_round_half_even(oracle):
    r = oracle.lib().sqzo_round0
    assert [r(0.5), r(1.5), r(2.5), r(-0.5), r(3.49999)] == [0.0, 2.0, 2.0, -0.0, 3.0]


def test_baf_bayes_independent_restatement(oracle):
    """A second, numpy/scipy restatement of R/bayes.R:61-79 must agree with the C one."""
    rng = np.random.default_rng(7)
    n = 200
    Bf = rng.uniform(0, 0.5, n)
    Bf[::17] = np.nan
    ratio = np.exp(rng.normal(0, 0.4, n))
    sdb, sdr = rng.uniform(0.03, 0.12, n), rng.uniform(0.3, 1.5, n)
    w = rng.uniform(3, 100, n)
    for (cell, plo, rp) in ((0.63, 2.7, False), (1.0, 3.1, False), (0.41, 1.3, True)):
        out = oracle.baf_bayes(Bf, ratio, cell, plo, 1.0, sdb, sdr, w, w, CNt_max=7,
                               priors_table={"CN": [2], "value": [2]}, ratio_priority=rp)
        ty = oracle.baf_types_matrix(0, 7)
        cnt, b = ty[:, 1], ty[:, 2]
        with np.errstate(divide="ignore", invalid="ignore"):
            mr = np.log(((1 - cell) + cnt / 2 * cell) / (plo / 2 * cell + 1 - cell))
            mbaf = (b * cell + (1 - cell)) / (cnt * cell + 2 * (1 - cell))
        pri = np.where(cnt == 2, 2.0, 1.0)
        pri = pri / pri.sum()
        for s in range(n):
            with np.errstate(invalid="ignore", divide="ignore"):
                sr = scipy_stats.t.logpdf((mr - np.log(ratio[s])) / (sdr[s] / np.sqrt(w[s])),
                                          5) + np.log(pri)
                post = sr if np.isnan(Bf[s]) else sr + scipy_stats.t.logpdf(
                    (mbaf - Bf[s]) / (sdb[s] / np.sqrt(w[s])), 5)
            if rp:
                cn = cnt[np.nanargmax(sr)]
                post = np.where(cnt == cn, post, np.nan)
            k = int(np.nanargmax(post))
            assert out[s, 0] == cnt[k]
            assert out[s, 3] == pytest.approx(post[k], rel=1e-12)
            if np.isnan(Bf[s]):
                assert np.isnan(out[s, 1]) and np.isnan(out[s, 2])
            else:
                assert out[s, 2] == b[k] and out[s, 1] == cnt[k] - b[k]


def test_mufreq_bayes_independent_restatement(oracle):
    rng = np.random.default_rng(11)
    n = 150
    F = rng.integers(0, 101, n) / 100.0
    ratio = np.exp(rng.normal(0, 0.4, n))
    wm = rng.integers(20, 300, n).astype(float)
    out = oracle.mufreq_bayes(F, ratio, 0.66, 2.3, 1.0, wm, 2 * wm, CNt_min=1, CNt_max=7,
                              priors_table={"CN": [2], "value": [2]})
    ty = oracle.mufreq_types_matrix(1, 7)
    cnt, mt = ty[:, 1], ty[:, 2]
    cell, plo = 0.66, 2.3
    mr = ((1 - cell) + cnt / 2 * cell) / (plo / 2 * cell + 1 - cell)
    mf = 1 - ((cnt - mt) * cell + 2 * (1 - cell)) / (cnt * cell + 2 * (1 - cell))
    mf[mf == 0] = 0.01
    pri = np.where(cnt == 2, 2.0, 1.0)
    pri = pri / pri.sum()
    for s in range(n):
        kr = np.rint(2 * wm[s] * (ratio[s] / (1 + ratio[s])))
        km = np.rint(F[s] * wm[s])
        sr = scipy_stats.binom.pmf(kr, 2 * wm[s], mr / (1 + mr)) * pri
        sm = scipy_stats.binom.pmf(km, wm[s], mf)
        post = sr * sm
        post[post == 0] = 1e-323
        cn = cnt[np.argmax(sr)]
        lp = np.where(cnt == cn, np.log(post), -np.inf)
        k = int(np.argmax(lp))
        assert out[s, 0] == 2 and out[s, 1] == cnt[k]
        assert out[s, 3] == pytest.approx(lp[k], rel=1e-9)
        # Mt can only differ from the scipy pick on a numerical tie
        assert out[s, 2] == mt[k] or abs(lp[k] - out[s, 3]) < 1e-9 * abs(lp[k])


def test_grid_sum_matches_point_sums(oracle):
    """baf_grid_sums (R/bayes.R:109-113) == long-double sum of baf_bayes()[, 4]."""
    rng = np.random.default_rng(3)
    n = 300
    Bf, ratio = rng.uniform(0, 0.5, n), np.exp(rng.normal(0, 0.3, n))
    pl, ce = np.array([1.5, 2.0, 3.3]), np.array([0.2, 0.9, 1.0])
    sums = oracle.baf_grid_sums(Bf, ratio, ce, pl, 1.0, 0.05, 0.8, 10, 10)
    for ic, c in enumerate(ce):
        for ip, p in enumerate(pl):
            o = oracle.baf_bayes(Bf, ratio, c, p, 1.0, 0.05, 0.8, 10, 10)
            ref = float(np.sum(o[:, 3].astype(np.longdouble)))
            assert sums[ic * 3 + ip] == pytest.approx(ref, rel=1e-15)
    fit = oracle.finish_model_fit(sums, pl, ce)
    assert fit["lpp"].shape == (3, 3) and abs(fit["lpp"].sum() - 1) < 1e-12
