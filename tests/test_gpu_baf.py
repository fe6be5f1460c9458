"""Parity of the CUDA baf path with the CPU oracle, through the C ABI
(sqz_baf_grid, sqz_baf_grid_batch, sqz_baf_point, sqz_model_points).

Bars (BASELINE.json): CNt / A / B and the selected (cellularity, ploidy) bit-exact; LPP
within 1e-9 relative in fp64 (per-segment LPP, per-grid-point sums, and the normalised
lpp matrix wherever it is above 1e-290)."""
import ctypes as C

import numpy as np
import pytest

import sequenza_b200 as S
from sequenza_b200 import _lib
from sequenza_b200.synth import baf_args, synth_segments

pytestmark = pytest.mark.gpu
RTOL = 1e-9
PRI = {"CN": [2], "value": [2]}  # sequenza.fit's default priors.table (R/fit.R:6)


def rel_close(a, b, rtol=RTOL):
    a, b = np.asarray(a), np.asarray(b)
    both_nan = np.isnan(a) & np.isnan(b)
    same_inf = np.isinf(a) & (a == b)
    with np.errstate(invalid="ignore"):
        ok = np.abs(a - b) <= rtol * np.abs(b)
    return bool(np.all(ok | both_nan | same_inf))


def check_point(oracle, seg, cell, plo, **kw):
    got = S.baf_bayes(cellularity=cell, ploidy=plo, **baf_args(seg), **kw)
    ref, margin = oracle.baf_bayes(cellularity=cell, ploidy=plo, return_margin=True,
                                   **baf_args(seg), **kw)
    # a state may differ only where the oracle's own top-2 gap is at rounding level;
    # exact ties (gap 0: e.g. Bf = NA makes every B of a CNt equal) must still match,
    # because both sides take the first maximum
    tie = (margin > 0) & (margin < 1e-12)
    for col in range(3):
        g, r = got[:, col], ref[:, col]
        same = (g == r) | (np.isnan(g) & np.isnan(r))
        assert np.all(same | tie), "column %d differs off-tie" % col
    assert tie.sum() <= 2
    assert rel_close(got[:, 3], ref[:, 3])
    # NA must be R's NA_real_ (payload 1954), not a plain NaN
    na = np.isnan(ref[:, 1])
    if na.any():
        assert set(got[na, 1].view(np.uint64).tolist()) == {0x7FF00000000007A2}
    return got, ref


def test_device_model_points_testthat(golden):
    """tests/testthat/test.model.R known answers, computed by the device code."""
    g = golden("testthat_model.json")

    def lookup(kind, cell, plo, cnt, second, cn_max=4):
        m = S.device_model_points(kind, cell, plo, 1.0, 0, cn_max, 2)
        k = np.flatnonzero((m["CNt"] == cnt) & (m["second"] == second))[0]
        return m["depth.ratio"][k], m["value"][k]

    for c in g["theoretical_depth_ratio"]["cases"]:
        assert lookup("baf", c["cellularity"], c["ploidy"], c["CNt"], 0)[0] == \
            pytest.approx(c["expect"], rel=1.5e-8)
    for c in g["theoretical_baf"]["cases"]:
        v = lookup("baf", c["cellularity"], 2.0, c["CNt"], c["B"])[1]
        assert (round(v, c["round"]) if c["round"] else v) == pytest.approx(c["expect"])
    for c in g["theoretical_mufreq"]["cases"]:
        v = lookup("mufreq", c["cellularity"], 2.0, c["CNt"], c["Mt"])[1]
        assert (round(v, c["round"]) if c["round"] else v) == pytest.approx(c["expect"], abs=1e-12)
    b = S.device_model_points("baf", 0.5, 2.0, 1.0, 0, 3, 2)
    assert b["CNt"].tolist() == g["baf_types_matrix"]["CNt"]
    assert b["second"].tolist() == g["baf_types_matrix"]["B"]
    m = S.device_model_points("mufreq", 0.5, 2.0, 1.0, 0, 3, 2)
    assert m["CNt"].tolist() == g["mufreq_types_matrix"]["CNt"]
    assert m["second"].tolist() == g["mufreq_types_matrix"]["Mt"]


def test_device_model_points_bitwise_vs_oracle(oracle):
    m = S.device_model_points("baf", 0.37, 2.9, 1.05, 0, 20, 2, PRI)
    for t in range(m["CNt"].size):
        assert m["depth.ratio"][t] == oracle.theoretical_depth_ratio(
            m["CNt"][t], 0.37, 2.9, 2, 2, 1.05)
        assert m["value"][t] == oracle.theoretical_baf(m["CNt"][t], m["second"][t], 0.37, 2)
    assert m["prior"].sum() == pytest.approx(1.0, rel=1e-15)
    assert m["prior"][m["CNt"] == 2][0] == 2.0 / 123.0


def test_vignette_rows_on_gpu(golden):
    """inst/doc/sequenza.html: six rows of baf.bayes output at the CP.example solution."""
    v = golden("vignette_rows.json")
    rows = np.array([[float(x) for x in r] for r in v["rows"]])
    t = {c: rows[:, i] for i, c in enumerate(v["columns"])}
    out = S.baf_bayes(Bf=t["Bf"], depth_ratio=t["depth.ratio"], cellularity=0.88, ploidy=2.0,
                      avg_depth_ratio=1.0, sd_Bf=t["sd.BAF"], sd_ratio=t["sd.ratio"],
                      weight_Bf=1, weight_ratio=(t["end.pos"] - t["start.pos"]) / 1e6,
                      CNt_max=20)
    assert out[:, 0].tolist() == t["CNt"].tolist()
    assert out[:, 1].tolist() == t["A"].tolist()
    assert out[:, 2].tolist() == t["B"].tolist()
    assert np.max(np.abs(out[:, 3] - t["LPP"])) < 5e-3


@pytest.mark.parametrize("cn_max,rp", [(7, False), (7, True), (20, False), (20, True)])
def test_baf_bayes_matches_oracle(oracle, cn_max, rp):
    seg = synth_segments(3000, cn_max=min(cn_max, 7), seed=101 + cn_max)
    for cell, plo in ((0.7, 2.4), (0.33, 4.1), (1.0, 2.0), (0.1, 1.0)):
        check_point(oracle, seg, cell, plo, CNt_max=cn_max, priors_table=PRI,
                    ratio_priority=rp)


def test_baf_bayes_scalars_na_and_cnn1(oracle):
    seg = synth_segments(500, seed=5)
    # scalar sd / weights recycle like data.frame() (R/bayes.R:43-45)
    a = S.baf_bayes(seg["Bf"], seg["depth_ratio"], 0.6, 2.2, 1.0)
    b = oracle.baf_bayes(seg["Bf"], seg["depth_ratio"], 0.6, 2.2, 1.0)
    assert np.array_equal(a[:, 0], b[:, 0]) and rel_close(a[:, 3], b[:, 3])
    # male X/Y call of sequenza.results (R/results.R:139-145): Bf = NA, CNn = 1
    na = np.full(500, np.nan)
    a = S.baf_bayes(na, seg["depth_ratio"], 0.6, 2.2, 1.0, sd_Bf=na, weight_Bf=na,
                    sd_ratio=seg["sd_ratio"], weight_ratio=seg["weight_ratio"], CNt_max=20,
                    CNn=1)
    b = oracle.baf_bayes(na, seg["depth_ratio"], 0.6, 2.2, 1.0, sd_Bf=na, weight_Bf=na,
                         sd_ratio=seg["sd_ratio"], weight_ratio=seg["weight_ratio"],
                         CNt_max=20, CNn=1)
    assert np.array_equal(a[:, 0], b[:, 0]) and rel_close(a[:, 3], b[:, 3])
    assert np.isnan(a[:, 1]).all() and np.isnan(a[:, 2]).all()


def test_baf_bayes_degenerate_rows(oracle):
    """Zero sd, zero / negative / NA depth ratio, cellularity 1: IEEE special values must
    take the same route as in R (dt(+-Inf) = -Inf, NaN skipped by which.max)."""
    Bf = np.array([0.3, 0.3, 0.3, np.nan, 0.25, 0.5])
    ratio = np.array([1.0, 0.0, -1.0, np.nan, 1.2, 1.0])
    sdb = np.array([0.05, 0.05, 0.05, 0.05, 0.0, 0.05])
    sdr = np.array([0.0, 0.5, 0.5, 0.5, 0.5, 0.5])
    for cell in (0.5, 1.0):
        a = S.baf_bayes(Bf, ratio, cell, 2.0, 1.0, sdb, sdr, 10, 10)
        b = oracle.baf_bayes(Bf, ratio, cell, 2.0, 1.0, sdb, sdr, 10, 10)
        for col in range(3):
            same = (a[:, col] == b[:, col]) | (np.isnan(a[:, col]) & np.isnan(b[:, col]))
            assert same.all(), (cell, col, a, b)
        assert rel_close(a[:, 3], b[:, 3])


def test_baf_model_view_lattice(oracle):
    """The baf.bayes call of baf.model.view (R/graphics.R:410-425): a 100 x 50 lattice
    expand.grid(Bf = seq(0.001, 0.5, 0.005), ratio = seq(0.01, 2.5, 0.05)), Bf fastest,
    scalar sd (the mean segment sd) and weights 10 / 10, at the fitted solution, with the
    CNt.max of sequenza.fit (7) and of sequenza.results (20, R/results.R:226)."""
    bf, ratio = S.r_seq(0.001, 0.5, 0.005), S.r_seq(0.01, 2.5, 0.05)
    assert bf.size == 100 and ratio.size == 50
    Bf = np.tile(bf, ratio.size)          # expand.grid: first factor varies fastest
    dr = np.repeat(ratio, bf.size)
    for cn_max in (7, 20):
        kw = dict(cellularity=0.88, ploidy=2.0, avg_depth_ratio=1.0, sd_Bf=0.0731,
                  weight_Bf=10, sd_ratio=0.6124, weight_ratio=10, CNt_max=cn_max, CNt_min=0,
                  CNn=2, ratio_priority=False)
        got = S.baf_bayes(Bf, dr, **kw)
        ref, margin = oracle.baf_bayes(Bf, dr, return_margin=True, **kw)
        tie = (margin > 0) & (margin < 1e-12)
        for col in range(3):
            assert np.all((got[:, col] == ref[:, col]) | tie), col
        assert rel_close(got[:, 3], ref[:, 3])
        z = got[:, 3].reshape(ratio.size, bf.size).T  # tapply(LPP, list(Bf, ratio), mean)
        assert z.shape == (100, 50) and np.isfinite(z).all()


def test_baf_bayes_empty():
    out = S.baf_bayes([], [], 0.5, 2.0, 1.0)
    assert out.shape == (0, 4)


@pytest.mark.parametrize("mode", [S.SQZ_MODE_FAST, S.SQZ_MODE_FAITHFUL])
@pytest.mark.parametrize("cn_max,rp", [(7, False), (7, True), (20, False)])
def test_grid_sums_match_oracle(oracle, mode, cn_max, rp):
    """Default grid of sequenza.fit (91 x 61, cellularity up to 1.0) on 1,500 segments."""
    seg = synth_segments(1500, seed=77)
    ce, pl = S.r_seq(0.1, 1, 0.01), S.r_seq(1, 7, 0.1)
    kw = dict(CNt_max=cn_max, priors_table=PRI, ratio_priority=rp)
    got = S.baf_grid_sums(cellularity=ce, ploidy=pl, mode=mode, **baf_args(seg), **kw)
    ref = oracle.baf_grid_sums(cellularity=ce, ploidy=pl, **baf_args(seg), **kw)
    assert rel_close(got, ref)
    assert int(np.argmax(got)) == int(np.argmax(ref))  # selected solution, bit-exact
    a, b = S.finish_model_fit(got, pl, ce), oracle.finish_model_fit(ref, pl, ce)
    big = b["lpp"] > 1e-290
    assert big.sum() > 10
    np.testing.assert_allclose(a["lpp"][big], b["lpp"][big], rtol=RTOL, atol=0)
    ca, cb = S.get_ci(a), oracle.get_ci(b)
    assert ca["max.ploidy"] == cb["max.ploidy"]
    assert ca["max.cellularity"] == cb["max.cellularity"]
    assert ca["confint.ploidy"] == tuple(cb["confint.ploidy"])
    assert ca["confint.cellularity"] == tuple(cb["confint.cellularity"])


def test_grid_ragged_sizes_and_chunks(oracle):
    """Grids that do not fill a ploidy tile, a single column, one segment, and a large N
    that is split into several segment chunks per CTA."""
    for n, ce, pl in ((1, [0.5], [2.0]), (33, [0.2, 0.9], S.r_seq(1, 4.3, 0.1)),
                      (257, S.r_seq(0.3, 0.4, 0.01), [1.0, 2.0, 3.0]),
                      (40000, [0.45, 0.7], S.r_seq(1.5, 3, 0.25))):
        seg = synth_segments(n, seed=n)
        got = S.baf_grid_sums(cellularity=ce, ploidy=pl, priors_table=PRI, **baf_args(seg))
        ref = oracle.baf_grid_sums(cellularity=ce, ploidy=pl, priors_table=PRI,
                                   **baf_args(seg))
        assert rel_close(got, ref), n


def test_grid_flagged_points_are_repaired(oracle):
    """Inputs the product kernel cannot represent (zero sd.ratio, zero depth ratio) must
    come back exactly as the faithful path computes them."""
    seg = synth_segments(400, seed=9)
    seg["sd_ratio"][7] = 0.0       # x = +-Inf -> dt = -Inf for every type
    ce, pl = [0.3, 0.8, 1.0], S.r_seq(1, 3, 0.5)
    got = S.baf_grid_sums(cellularity=ce, ploidy=pl, priors_table=PRI, **baf_args(seg))
    ref = oracle.baf_grid_sums(cellularity=ce, ploidy=pl, priors_table=PRI, **baf_args(seg))
    assert np.isneginf(ref).all() and np.array_equal(got, ref)
    seg = synth_segments(400, seed=10)
    seg["depth_ratio"][3] = np.nan  # every posterior NaN: the reference stops; we give NaN
    got = S.baf_grid_sums(cellularity=ce, ploidy=pl, priors_table=PRI, **baf_args(seg))
    assert np.isnan(got).all()
    seg = synth_segments(400, seed=11)
    seg["sd_Bf"][:] = 1e-200        # products overflow: flagged, then repaired
    got = S.baf_grid_sums(cellularity=ce, ploidy=pl, priors_table=PRI, **baf_args(seg))
    ref = oracle.baf_grid_sums(cellularity=ce, ploidy=pl, priors_table=PRI, **baf_args(seg))
    assert rel_close(got, ref)


def test_grid_empty_inputs():
    assert S.baf_grid_sums([], [], [0.5], [2.0], 1.0).tolist() == [0.0]
    assert S.baf_grid_sums([0.3], [1.0], [], [2.0], 1.0).size == 0


def test_grid_batch_equals_single(oracle):
    """Cohort launch (BASELINE config 5): ragged samples, one call; identical to fitting
    each sample on its own."""
    ce, pl = S.r_seq(0.2, 1, 0.05), S.r_seq(1, 5, 0.25)
    samples = []
    for k, n in enumerate((120, 1, 333, 64, 2000)):
        seg = synth_segments(n, cellularity=0.3 + 0.1 * k, ploidy=1.8 + 0.4 * k, seed=40 + k,
                             avg_depth_ratio=1.0 + 0.01 * k)
        samples.append(baf_args(seg))
    fits = S.baf_model_fit_batch(samples, cellularity=ce, ploidy=pl, priors_table=PRI)
    assert len(fits) == 5
    # the same cohort handed over as concatenated vectors + offsets
    keys = ("Bf", "depth_ratio", "sd_Bf", "sd_ratio", "weight_Bf", "weight_ratio")
    cohort = {k: np.concatenate([np.asarray(s[k], dtype=float) for s in samples]) for k in keys}
    cohort["seg_offsets"] = np.cumsum([0] + [len(s["depth_ratio"]) for s in samples])
    cohort["avg_depth_ratio"] = [s["avg_depth_ratio"] for s in samples]
    for f, g in zip(fits, S.baf_model_fit_batch(cohort, cellularity=ce, ploidy=pl,
                                                priors_table=PRI)):
        np.testing.assert_array_equal(f["lpp"], g["lpp"])
    for smp, f in zip(samples, fits):
        one = S.baf_model_fit(cellularity=ce, ploidy=pl, priors_table=PRI, **smp)
        np.testing.assert_array_equal(f["lpp"], one["lpp"])
        ref = oracle.baf_model_fit(cellularity=ce, ploidy=pl, priors_table=PRI, **smp)
        big = ref["lpp"] > 1e-290
        np.testing.assert_allclose(f["lpp"][big], ref["lpp"][big], rtol=RTOL)


def test_sequenza_fit_end_to_end(oracle):
    """sequenza.fit -> get.ci on a synthetic extract: the reported solution equals the
    oracle's, and lies at the simulated truth."""
    seg = synth_segments(600, cellularity=0.62, ploidy=3.1, seed=12)
    n = seg["Bf"].size
    length = seg["weight_ratio"] * 1e6
    tab = {"chromosome": np.array(["1"] * n), "start.pos": np.zeros(n), "end.pos": length,
           "Bf": seg["Bf"], "N.BAF": np.full(n, 50.0), "sd.BAF": seg["sd_Bf"],
           "depth.ratio": seg["depth_ratio"], "N.ratio": np.full(n, 80.0),
           "sd.ratio": seg["sd_ratio"]}
    cp = S.sequenza_fit({"segments": [tab], "avg.depth.ratio": 1.0})
    ref = oracle.baf_model_fit(cellularity=S.r_seq(0.1, 1, 0.01), ploidy=S.r_seq(1, 7, 0.1),
                               priors_table=PRI, **baf_args(seg))
    ci, ci_ref = S.get_ci(cp), oracle.get_ci(ref)
    assert ci["max.cellularity"] == ci_ref["max.cellularity"]
    assert ci["max.ploidy"] == ci_ref["max.ploidy"]
    assert abs(ci["max.cellularity"] - 0.62) <= 0.02 and abs(ci["max.ploidy"] - 3.1) <= 0.1


def test_full_size_properties_config2():
    """BASELINE config 2 at full size (3e4 segments, CNt.max = 20, default grid): the
    oracle would need minutes, so check size-independent properties instead."""
    seg = synth_segments(30000, seed=S.synth.DEFAULT_SEED)
    ce, pl = S.r_seq(0.1, 1, 0.01), S.r_seq(1, 7, 0.1)
    kw = dict(cellularity=ce, ploidy=pl, CNt_max=20, priors_table=PRI)
    s1 = S.baf_grid_sums(**baf_args(seg), **kw)
    assert np.isfinite(s1).all()
    # (1) a permutation of the segments changes nothing beyond summation order
    perm = np.random.default_rng(0).permutation(30000)
    seg2 = {k: (v[perm] if isinstance(v, np.ndarray) else v) for k, v in baf_args(seg).items()}
    s2 = S.baf_grid_sums(**seg2, **kw)
    np.testing.assert_allclose(s2, s1, rtol=1e-13)
    # (2) additivity over a split of the segment table
    h = {k: (v[:12345] if isinstance(v, np.ndarray) else v) for k, v in baf_args(seg).items()}
    t = {k: (v[12345:] if isinstance(v, np.ndarray) else v) for k, v in baf_args(seg).items()}
    np.testing.assert_allclose(S.baf_grid_sums(**h, **kw) + S.baf_grid_sums(**t, **kw), s1,
                               rtol=1e-13)
    # (3) fast and faithful kernels agree on a column sample of the same full-size input
    cols = ce[[0, 30, 60, 90]]
    kw2 = dict(kw, cellularity=cols)
    np.testing.assert_allclose(
        S.baf_grid_sums(**baf_args(seg), **kw2),
        S.baf_grid_sums(mode=S.SQZ_MODE_FAITHFUL, **baf_args(seg), **kw2), rtol=1e-12)
    # (4) normalised, peaked at the simulated truth, and idempotent
    fit = S.finish_model_fit(s1, pl, ce)
    assert abs(fit["lpp"].sum() - 1) < 1e-9
    ci = S.get_ci(fit)
    assert abs(ci["max.cellularity"] - 0.7) <= 0.011 and abs(ci["max.ploidy"] - 2.4) <= 0.11
    np.testing.assert_array_equal(S.baf_grid_sums(**baf_args(seg), **kw), s1)


def test_grid_kernel_paths_agree(oracle, monkeypatch):
    """The three fast-path kernels (fused, split, split + candidate pruning) and the
    orderings of the ploidy axis that switch pruning off must all give the oracle's sums."""
    seg = synth_segments(3000, seed=88)
    ce, pl = S.r_seq(0.1, 1, 0.02), S.r_seq(1, 7, 0.05)
    kw = dict(cellularity=ce, priors_table=PRI, **baf_args(seg))
    ref = oracle.baf_grid_sums(ploidy=pl, **kw)
    for env in ({"SQZ_SPLIT": "0"}, {"SQZ_PRUNE": "0"}, {"SQZ_PRUNE": "1"},
                {"SQZ_PRUNE": "1", "SQZ_PT": "16"}, {"SQZ_PRUNE": "1", "SQZ_PT": "8"},
                {"SQZ_COLWARP": "0"}, {"SQZ_COLWARP": "0", "SQZ_PT": "16"},
                {"SQZ_COLWARP": "1", "SQZ_PT": "32"}, {"SQZ_COLWARP": "1", "SQZ_PT": "16"}):
        for k in ("SQZ_SPLIT", "SQZ_PRUNE", "SQZ_PT", "SQZ_COLWARP"):
            monkeypatch.delenv(k, raising=False)
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        got = S.baf_grid_sums(ploidy=pl, **kw)
        assert rel_close(got, ref), env
        assert int(np.argmax(got)) == int(np.argmax(ref)), env
    for k in ("SQZ_SPLIT", "SQZ_PRUNE", "SQZ_PT", "SQZ_COLWARP"):
        monkeypatch.delenv(k, raising=False)
    # descending and unsorted ploidy vectors (pruning needs a monotone tile; unsorted input
    # falls back to the plain split kernel), duplicated values
    for p2 in (pl[::-1].copy(), np.random.default_rng(1).permutation(pl),
               np.concatenate([pl[:5], pl[:5]])):
        got = S.baf_grid_sums(ploidy=p2, **kw)
        assert rel_close(got, oracle.baf_grid_sums(ploidy=p2, **kw))
    # CNt.max = 20 with pruning forced on (many copy numbers, per-thread candidates)
    monkeypatch.setenv("SQZ_PRUNE", "1")
    kw20 = dict(kw, CNt_max=20)
    assert rel_close(S.baf_grid_sums(ploidy=pl, **kw20), oracle.baf_grid_sums(ploidy=pl, **kw20))


@pytest.mark.parametrize("cn_max", [7, 20])
@pytest.mark.parametrize("n_seg", [1, 3, 4, 257, 2050])
@pytest.mark.parametrize("grid", [(46, 121), (33, 40), (5, 20), (91, 5), (64, 601)])
def test_grid_column_lane_kernel(oracle, monkeypatch, n_seg, grid, cn_max):
    """Column-per-lane pruned kernel (sqz_baf_colwarp.cu), forced on: ragged column groups,
    fewer than four ploidy tiles per CTA, ragged last tile, segment counts that are not a
    multiple of the stage size, several chunks; and the same call as a cohort."""
    monkeypatch.setenv("SQZ_COLWARP", "1")
    nC, nP = grid
    if n_seg * nC * nP * (5 if cn_max == 20 else 1) > 2e7:
        pytest.skip("oracle time")
    ce = np.linspace(0.05, 1.0, nC)
    pl = np.linspace(1.0, 7.0, nP)
    seg = synth_segments(n_seg, seed=300 + n_seg, cn_max=cn_max)
    kw = dict(cellularity=ce, ploidy=pl, priors_table=PRI, CNt_max=cn_max)
    ref = oracle.baf_grid_sums(**baf_args(seg), **kw)
    got = S.baf_grid_sums(**baf_args(seg), **kw)
    assert rel_close(got, ref)
    assert int(np.nanargmax(got)) == int(np.nanargmax(ref))
    if n_seg == 2050 and nP <= 121:
        monkeypatch.setenv("SQZ_CHUNKS", "7")
        assert rel_close(S.baf_grid_sums(**baf_args(seg), **kw), ref)
        monkeypatch.delenv("SQZ_CHUNKS")
        other = synth_segments(77, seed=9, cellularity=0.4)
        fits = S.baf_model_fit_batch([baf_args(seg), baf_args(other), baf_args(seg)], ce, pl,
                                     priors_table=PRI, n_gpus=1, CNt_max=cn_max)
        one = S.baf_model_fit(cellularity=ce, ploidy=pl, priors_table=PRI, n_gpus=1,
                              CNt_max=cn_max, **baf_args(other))
        np.testing.assert_allclose(fits[1]["lpp"], one["lpp"], rtol=1e-9, atol=1e-300)
        np.testing.assert_array_equal(fits[0]["lpp"], fits[2]["lpp"])


@pytest.mark.parametrize("scale", [1e-2, 1e-5, 1e-9, 1e-21, 1e3])
def test_column_lane_kernel_fp32_bounds_at_extreme_scales(oracle, monkeypatch, scale):
    """The column-per-lane kernel for CNt 0..7 takes its candidate bounds in fp32, rounded
    downwards (sqz_baf_colwarp.cu).  Standard deviations scaled far down make the
    standardised residuals, the BAF factors and the first candidate's maximum leave the
    range (and the precision) of a float; scaled up, every copy number is a candidate.  A
    bound that is too high would drop the true minimum: the sums must still equal the
    oracle's, or the grid point must have been handed to the faithful kernel."""
    monkeypatch.setenv("SQZ_COLWARP", "1")
    ce, pl = np.linspace(0.1, 1.0, 64), S.r_seq(1, 7, 0.05)
    seg = synth_segments(600, seed=77)
    seg["sd_ratio"] = seg["sd_ratio"] * scale
    seg["sd_Bf"] = seg["sd_Bf"] * (scale if scale > 1e-12 else scale * 1e-20)
    # a few segments that sit exactly on a model value, and some far away from all of them
    seg["depth_ratio"][:5] = [0.5, 1.0, 1.5, 2.0, 1e-3]
    seg["depth_ratio"][5:8] = [37.0, 1e4, 4e-8]
    kw = dict(cellularity=ce, ploidy=pl, priors_table=PRI)
    L = _lib.lib()
    L.sqz_profile_enable(1)
    try:
        got = S.baf_grid_sums(**baf_args(seg), **kw)
        path, n_eval, launches = C.c_int(-1), C.c_ulonglong(0), C.c_longlong(0)
        L.sqz_profile_counters(C.byref(path), C.byref(n_eval), C.byref(launches))
    finally:
        L.sqz_profile_enable(0)
    assert path.value == 3
    ref = oracle.baf_grid_sums(**baf_args(seg), **kw)
    fin = np.isfinite(ref)
    assert np.array_equal(np.isfinite(got), fin)
    assert fin.any()
    assert np.max(np.abs(got[fin] - ref[fin]) / np.abs(ref[fin])) <= 1e-9
    assert int(np.nanargmax(got)) == int(np.nanargmax(ref))


def test_full_size_properties_config3():
    """BASELINE config 3 at full size (1e5 segments, CNt.max = 7, ploidy step 0.01: the
    pruned kernel): fast == faithful on a sample of full columns, additivity over a split
    of the segment table, and the fine-grid solution sits at the simulated truth."""
    seg = synth_segments(100000, seed=S.synth.DEFAULT_SEED)
    pl = S.r_seq(1, 7, 0.01)
    cols = np.array([0.1, 0.463, 0.7, 0.999, 1.0])
    kw = dict(cellularity=cols, ploidy=pl, priors_table=PRI)
    fast = S.baf_grid_sums(**baf_args(seg), **kw)
    slow = S.baf_grid_sums(mode=S.SQZ_MODE_FAITHFUL, **baf_args(seg), **kw)
    assert np.isfinite(slow).all()
    np.testing.assert_allclose(fast, slow, rtol=1e-12, atol=0)
    assert int(np.argmax(fast)) == int(np.argmax(slow))
    h = {k: (v[:33333] if isinstance(v, np.ndarray) else v) for k, v in baf_args(seg).items()}
    t = {k: (v[33333:] if isinstance(v, np.ndarray) else v) for k, v in baf_args(seg).items()}
    np.testing.assert_allclose(S.baf_grid_sums(**h, **kw) + S.baf_grid_sums(**t, **kw), fast,
                               rtol=1e-13)
    ce = S.r_seq(0.6, 0.8, 0.001)
    fit = S.baf_model_fit(cellularity=ce, ploidy=pl, priors_table=PRI, **baf_args(seg))
    ci = S.get_ci(fit)
    assert abs(ci["max.cellularity"] - 0.7) <= 0.005 and abs(ci["max.ploidy"] - 2.4) <= 0.02
    # normalised lpp: device finish (fixed-point masses) against the host steps
    host = S.finish_model_fit(S.baf_grid_sums(cellularity=ce, ploidy=pl, priors_table=PRI,
                                              **baf_args(seg)), pl, ce, device=False)
    np.testing.assert_allclose(fit["lpp"], host["lpp"], rtol=1e-9, atol=1e-300)
    assert S.get_ci(host)["confint.ploidy"] == ci["confint.ploidy"]
