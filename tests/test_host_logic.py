"""Host-side mirror of the reference interface (sequenza_b200/{model,bayes,fit,
solutions}.py): everything that stays R in the reference.  CPU only."""
import numpy as np
import pytest

import sequenza_b200 as S
from sequenza_b200 import bayes, fit as fit_mod


def test_r_seq_matches_reference_defaults():
    # R/fit.R:4: seq(0.1, 1, 0.01) has 91 values, seq(1, 7, 0.1) has 61
    c, p = S.r_seq(0.1, 1, 0.01), S.r_seq(1, 7, 0.1)
    assert c.size == 91 and p.size == 61
    assert c[0] == 0.1 and c[-1] == 1.0 and p[-1] == 7.0
    assert c[78] == 0.1 + 78 * 0.01  # from + k*by, not a running sum
    assert S.r_seq(0.3, 1, 0.01).size == 71  # R/bayes.R:103


def test_model_helpers_testthat(golden):
    g = golden("testthat_model.json")
    for c in g["theoretical_depth_ratio"]["cases"]:
        assert float(S.theoretical_depth_ratio(c["CNt"], c["cellularity"], c["ploidy"])) == \
            pytest.approx(c["expect"], rel=1.5e-8)
    for c in g["theoretical_baf"]["cases"]:
        v = float(S.theoretical_baf(c["CNt"], c["B"], c["cellularity"]))
        assert (round(v, c["round"]) if c["round"] else v) == pytest.approx(c["expect"])
    for c in g["theoretical_mufreq"]["cases"]:
        v = float(S.theoretical_mufreq(c["CNt"], c["Mt"], c["cellularity"]))
        assert (round(v, c["round"]) if c["round"] else v) == pytest.approx(c["expect"], abs=1e-12)
    b = S.baf_types_matrix(**g["baf_types_matrix"]["args"])
    assert b["CNt"].tolist() == g["baf_types_matrix"]["CNt"]
    assert b["B"].tolist() == g["baf_types_matrix"]["B"]
    m = S.mufreq_types_matrix(**g["mufreq_types_matrix"]["args"])
    assert m["CNt"].tolist() == g["mufreq_types_matrix"]["CNt"]
    assert m["Mt"].tolist() == g["mufreq_types_matrix"]["Mt"]
    assert np.isnan(S.theoretical_baf(2, 1, 0.5, CNn=1))  # R/model.R:12


def test_model_points_match_oracle(oracle):
    ty = S.baf_types_matrix(0, 7)
    mp = S.baf_model_points(0.37, 2.9, ty, 1.05)
    for t in range(ty["CNt"].size):
        assert mp["depth.ratio"][t] == oracle.theoretical_depth_ratio(
            ty["CNt"][t], 0.37, 2.9, 2, 2, 1.05)
        assert mp["BAF"][t] == oracle.theoretical_baf(ty["CNt"][t], ty["B"][t], 0.37, 2)
    tm = S.mufreq_types_matrix(1, 7)
    mm = S.mufreq_model_points(0.37, 2.9, tm, 1.05)
    for t in range(tm["CNt"].size):
        assert mm["mufreqs"][t] == oracle.theoretical_mufreq(tm["CNt"][t], tm["Mt"][t], 0.37, 2)


def test_finish_model_fit_matches_oracle(oracle):
    rng = np.random.default_rng(5)
    pl, ce = S.r_seq(1, 7, 0.1), S.r_seq(0.1, 1, 0.01)
    sums = -7e3 + 40 * rng.standard_normal(pl.size * ce.size)
    a, b = S.finish_model_fit(sums, pl, ce), oracle.finish_model_fit(sums, pl, ce)
    assert a["lpp"].shape == (61, 91)
    np.testing.assert_array_equal(a["ploidy"], b["ploidy"])
    np.testing.assert_array_equal(a["cellularity"], b["cellularity"])
    np.testing.assert_allclose(a["lpp"], b["lpp"], rtol=1e-13, atol=0)
    assert abs(a["lpp"].sum() - 1) < 1e-12
    # layout: rows = ploidy, columns = cellularity; input is ploidy-fastest
    k = int(np.argmax(sums))
    ip, ic = np.unravel_index(np.argmax(a["lpp"]), a["lpp"].shape)
    assert (ic, ip) == (k // pl.size, k % pl.size)
    # axis values went through 15-significant-digit labels (R/bayes.R:118-119)
    assert a["cellularity"].tolist() == [float("%.15g" % v) for v in ce]
    assert a["cellularity"][78] == 0.88


def test_finish_model_fit_duplicates_and_nan(oracle):
    pl, ce = np.array([2.0, 1.0, 2.0]), np.array([0.5, 0.25])
    sums = np.array([-10.0, -11.0, -12.0, -9.0, -13.0, -9.5])
    a, b = S.finish_model_fit(sums, pl, ce), oracle.finish_model_fit(sums, pl, ce)
    assert a["ploidy"].tolist() == [1.0, 2.0] and a["cellularity"].tolist() == [0.25, 0.5]
    np.testing.assert_allclose(a["lpp"], b["lpp"], rtol=1e-14)
    s2 = sums.copy()
    s2[1] = np.nan  # sum() has no na.rm (R/bayes.R:121): everything becomes NaN
    assert np.isnan(S.finish_model_fit(s2, pl, ce)["lpp"]).all()


def test_get_ci_matches_oracle_and_golden(oracle, golden):
    d = golden("cp_example.json")
    nr, nc = d["lpp_dim"]
    cp = {"ploidy": np.array(d["ploidy"]), "cellularity": np.array(d["cellularity"]),
          "lpp": np.array(d["lpp_column_major"]).reshape(nc, nr).T}
    a, b = S.get_ci(cp), oracle.get_ci(cp)
    assert a["max.ploidy"] == b["max.ploidy"] == 2.0
    assert a["max.cellularity"] == b["max.cellularity"] == 0.88
    assert a["confint.ploidy"] == tuple(b["confint.ploidy"])
    assert a["confint.cellularity"] == tuple(b["confint.cellularity"])
    np.testing.assert_allclose(a["values.ploidy"]["y"], b["values.ploidy"]["y"], rtol=1e-14)
    np.testing.assert_allclose(a["values.cellularity"]["x"], b["values.cellularity"]["x"],
                               rtol=1e-14)
    for lvl in (0.5, 0.999):
        x, y = S.get_ci(cp, lvl), oracle.get_ci(cp, lvl)
        assert x["confint.ploidy"] == tuple(y["confint.ploidy"])
        assert x["confint.cellularity"] == tuple(y["confint.cellularity"])
    alt = S.alternative_cp_solutions(cp)
    assert alt["SLPP"][0] == cp["lpp"].max()
    assert alt["ploidy"][0] == 2.0 and alt["cellularity"][0] == 0.88
    assert np.all(np.diff(alt["SLPP"]) <= 0)


def _extract():
    seg1 = {"chromosome": ["1"] * 4, "start.pos": [0, 5e6, 2e7, 3e7],
            "end.pos": [4e6, 9e6, 2.2e7, 9e7], "Bf": [0.4, 0.3, 0.2, np.nan],
            "N.BAF": [50, 1, 40, 30], "sd.BAF": [0.05, 0.06, 0.0, 0.08],
            "depth.ratio": [1.0, 1.2, 0.8, 1.5], "N.ratio": [100, 100, 5, 80],
            "sd.ratio": [0.5, 0.0, 0.7, 0.9]}
    segx = {"chromosome": ["X"], "start.pos": [0], "end.pos": [5e7], "Bf": [0.1],
            "N.BAF": [90], "sd.BAF": [0.04], "depth.ratio": [0.6], "N.ratio": [200],
            "sd.ratio": [0.4]}
    segy = dict(segx, chromosome=["Y"])
    return {"segments": [seg1, segx, segy], "avg.depth.ratio": 1.03,
            "mutations": [
                {"chromosome": ["1", "1", "1"], "F": [0.05, 0.3, 0.5],
                 "good.reads": [40.5, 33.5, np.nan], "adjusted.ratio": [1.0, 1.1, 0.9]},
                {"chromosome": ["X"], "F": [0.4], "good.reads": [21.0],
                 "adjusted.ratio": [0.7]}]}


def test_sequenza_fit_filters_and_dispatch(monkeypatch):
    seen = {}

    def fake_baf(**kw):
        seen.update(kw)
        return "baf"

    monkeypatch.setattr(fit_mod, "baf_model_fit", fake_baf)
    assert S.sequenza_fit(_extract()) == "baf"
    # kept: seg1 rows 0 (4 Mb) and 3 (60 Mb, Bf NA) and X; row 1 fails N.BAF > 1, row 2
    # fails N.ratio > 10 and the length filter; Y dropped for female (R/fit.R:34-42)
    np.testing.assert_array_equal(seen["depth_ratio"], [1.0, 1.5, 0.6])
    np.testing.assert_array_equal(seen["weight_ratio"], [4.0, 60.0, 50.0])
    np.testing.assert_array_equal(seen["weight_Bf"], seen["weight_ratio"])
    np.testing.assert_array_equal(seen["sd_ratio"], [0.5, 0.9, 0.4])
    assert np.isnan(seen["Bf"][1])
    assert seen["avg_depth_ratio"] == 1.03
    assert seen["priors_table"] == {"CN": [2], "value": [2]}  # R/fit.R:6
    assert seen["cellularity"].size == 91 and seen["ploidy"].size == 61
    assert seen["ratio_priority"] is False
    S.sequenza_fit(_extract(), female=False)
    np.testing.assert_array_equal(seen["depth_ratio"], [1.0, 1.5])  # X dropped too
    S.sequenza_fit(_extract(), segment_filter=0, N_ratio_filter=1, N_BAF_filter=0)
    # zero sd replaced by the column maximum before filtering (R/fit.R:29-32)
    np.testing.assert_array_equal(seen["sd_ratio"], [0.5, 0.9, 0.7, 0.9, 0.4])
    np.testing.assert_array_equal(seen["sd_Bf"], [0.05, 0.06, 0.08, 0.08, 0.04])
    S.sequenza_fit(_extract(), chromosome_list=[2])
    np.testing.assert_array_equal(seen["depth_ratio"], [0.6])


def test_sequenza_fit_mufreq_and_errors(monkeypatch):
    seen = {}
    monkeypatch.setattr(fit_mod, "mufreq_model_fit", lambda **kw: seen.update(kw) or "muf")
    assert S.sequenza_fit(_extract(), method="mufreq") == "muf"
    # F >= 0.1, NA rows excluded, w = round(good.reads) half-even, weight.ratio = 2 w
    np.testing.assert_array_equal(seen["mufreq"], [0.3, 0.4])
    np.testing.assert_array_equal(seen["weight_mufreq"], [34.0, 21.0])
    np.testing.assert_array_equal(seen["weight_ratio"], [68.0, 42.0])
    assert "ratio_priority" not in seen  # not forwarded on this path (R/fit.R:61-65)
    with pytest.raises(ValueError, match="only available methods"):
        S.sequenza_fit(_extract(), method="nope")


def test_priors_and_recycling():
    cn, v = bayes._priors(None, 0, 7)
    assert cn.tolist() == list(range(8)) and v.tolist() == [1.0] * 8
    cn, v = bayes._priors({"CN": [2], "value": [2]}, 0, 7)
    assert cn.tolist() == [2.0] and v.tolist() == [2.0]
    from sequenza_b200 import _lib
    assert _lib.vec(3, 4).tolist() == [3.0] * 4
    with pytest.raises(ValueError):
        _lib.vec([1, 2], 3)


def test_synth_is_seeded():
    from sequenza_b200.synth import synth_mutations, synth_segments
    a, b = synth_segments(500), synth_segments(500)
    np.testing.assert_array_equal(a["depth_ratio"], b["depth_ratio"])
    assert 0 < np.isnan(a["Bf"]).sum() < 30
    assert (a["depth_ratio"] > 0).all() and np.nanmax(a["Bf"]) <= 0.5
    m = synth_mutations(500)
    assert m["mufreq"].min() >= 0 and m["mufreq"].max() <= 1
