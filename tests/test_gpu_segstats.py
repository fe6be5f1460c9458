"""sqz_segment_stats (segment.breaks' per-segment statistics, R/segments.R:113-198, and the
b_allele_freq density mode, R/model.R:69-98) against the CPU oracle, through the C ABI.
Counts bit-exact; means, standard deviations and the BAF mode within 1e-12 relative (the
sums are double-double here and long double in the oracle; the mode is a grid point of
density()'s 512-point axis, whose bandwidth goes through pow())."""
import numpy as np
import pytest

import sequenza_b200 as S
from sequenza_b200 import segments as G

pytestmark = pytest.mark.gpu
RTOL = 1e-12


def both(oracle, g, min_reads=1, weighted=True):
    args = (g["row_off"], g["seg_off"], g["position"], g["adjusted_ratio"], g["depth_normal"],
            g["good_reads"], g["Af"], g["Bf"], g["is_het"], g["breaks_vect"], min_reads, weighted)
    return G.segment_stats(*args), oracle.segment_stats(*args)


def check(got, ref):
    assert got.shape == ref.shape
    np.testing.assert_array_equal(got[:, [1, 4]], ref[:, [1, 4]])  # N.BAF, N.ratio
    for col in (0, 2, 3, 5):
        a, b = got[:, col], ref[:, col]
        assert np.array_equal(np.isnan(a), np.isnan(b)), (col, a, b)
        ok = ~np.isnan(a)
        np.testing.assert_allclose(a[ok], b[ok], rtol=RTOL, atol=0, err_msg="column %d" % col)
    # NA (not a plain NaN) where the reference has NA: sd of < 2 values, Bf without het rows
    na = np.isnan(ref[:, 2]) & (ref[:, 1] < 2)
    if na.any():
        assert set(got[na, 2].view(np.uint64).tolist()) == {0x7FF00000000007A2}


@pytest.mark.parametrize("weighted", [True, False])
@pytest.mark.parametrize("shape", [(3, 4000, 7), (1, 300, 1), (5, 20000, 40), (2, 60, 9)])
def test_segment_stats_match_oracle(oracle, shape, weighted):
    g = G.synth_seqz(n_chr=shape[0], rows_per_chr=shape[1], segs_per_chr=shape[2],
                     seed=sum(shape))
    got, ref = both(oracle, g, weighted=weighted)
    check(got, ref)
    assert (ref[:, 4] > 0).any()


def test_segment_stats_edge_cases(oracle):
    g = G.synth_seqz(n_chr=1, rows_per_chr=500, segs_per_chr=5, seed=3)
    s0 = slice(0, 500)
    brk = g["breaks_vect"]
    pos = g["position"]
    seg_rows = [np.flatnonzero((pos > brk[2 * k]) & (pos <= brk[2 * k + 1])) for k in range(5)]
    g["is_het"][seg_rows[0]] = 0                                  # no het row: Bf NA, sd NA
    g["is_het"][seg_rows[1]] = 0
    g["is_het"][seg_rows[1][3]] = 1                               # one het row: Bf itself
    g["is_het"][seg_rows[2]] = 0
    g["is_het"][seg_rows[2][:2]] = 1                              # two identical: sd exactly 0
    g["Bf"][seg_rows[2][:2]] = 0.25
    g["Af"][seg_rows[2][:2]] = 0.75
    g["good_reads"][seg_rows[2][:2]] = 40.0
    g["good_reads"][seg_rows[3][::2]] = 0.0                       # below min.reads.baf
    g["adjusted_ratio"][seg_rows[4]] = np.nan                     # every ratio NA
    g["depth_normal"][seg_rows[3][0]] = 0.0                       # w.r = 0: 0/0 dropped by na.rm
    got, ref = both(oracle, g, min_reads=1)
    check(got, ref)
    assert np.isnan(got[0, 0]) and got[0, 1] == 0
    assert got[1, 0] == g["Bf"][seg_rows[1][3]] and np.isnan(got[1, 2])
    assert got[2, 2] == 0.0 and got[2, 1] == 2
    # all ratios NA: sum(rw, na.rm) / sum(w.r, na.rm) = 0 / positive, sd of nothing is NA
    assert got[4, 3] == 0.0 and np.isnan(got[4, 5]) and got[4, 4] == seg_rows[4].size
    # cut(): a row AT a segment's start is outside it, one AT its end inside
    k = 2
    assert pos[seg_rows[k][-1]] <= brk[2 * k + 1] and (pos == brk[2 * k]).any()
    assert got[k, 4] == seg_rows[k].size
    got, ref = both(oracle, g, min_reads=30, weighted=False)
    check(got, ref)
    del s0


def test_segment_stats_empty_and_errors():
    e = np.empty(0)
    out = G.segment_stats([0], [0], e, e, e, e, e, e, np.empty(0, dtype=np.uint8), e)
    assert out.shape == (0, 6)
    g = G.synth_seqz(n_chr=1, rows_per_chr=50, segs_per_chr=2, seed=1)
    g["position"][10], g["position"][11] = g["position"][11], g["position"][10]
    with pytest.raises(ValueError, match="sorted"):
        G.segment_stats(g["row_off"], g["seg_off"], g["position"], g["adjusted_ratio"],
                        g["depth_normal"], g["good_reads"], g["Af"], g["Bf"], g["is_het"],
                        g["breaks_vect"])


def test_segment_breaks_host_mirror(oracle):
    """segment.breaks end to end: unsorted rows of two chromosomes, duplicated break positions
    (unique.breaks), chromosome order of first appearance, the N.ratio per Mb filter."""
    rng = np.random.default_rng(4)
    n = 3000
    chrom = np.array(["chr2"] * n + ["chr1"] * n)
    pos = np.concatenate([rng.permutation(np.arange(1, 4 * n, 4)),
                          rng.permutation(np.arange(1, 4 * n, 4))]).astype(float)
    gr = np.round(rng.uniform(20, 80, 2 * n))
    b = rng.binomial(gr.astype(int), 0.3) / gr
    b = np.minimum(b, 1 - b)
    tab = {"chromosome": chrom, "position": pos,
           "zygosity.normal": np.where(rng.random(2 * n) < 0.4, "het", "hom"),
           "good.reads": gr, "Af": 1 - b, "Bf": b,
           "adjusted.ratio": np.exp(rng.normal(0, 0.3, 2 * n)),
           "depth.normal": np.round(rng.uniform(10, 60, 2 * n))}
    breaks = {"chrom": np.array(["chr2", "chr2", "chr1", "chr1", "chr1"]),
              "start.pos": np.array([1.0, 5000.0, 1.0, 3001.0, 9000.0]),
              "end.pos": np.array([5000.0, 11999.0, 3000.0, 9000.0, 9010.0])}
    out = G.segment_breaks(tab, breaks)
    assert list(out["chromosome"][:2]) == ["chr2", "chr2"]
    assert out["start.pos"][1] == 5001.0          # unique.breaks bumped the repeated 5000
    assert (out["N.ratio"] / ((out["end.pos"] - out["start.pos"]) / 1e6) >= 2).all()
    # against the oracle on the same layout
    order, names, ro, so, brk = G.layout_genome(chrom, pos, breaks)
    ref = oracle.segment_stats(ro, so, pos[order], tab["adjusted.ratio"][order],
                               tab["depth.normal"][order], gr[order], tab["Af"][order],
                               b[order], (tab["zygosity.normal"][order] == "het"), brk)
    keep = (ref[:, 4] / ((brk[1::2] - brk[0::2]) / 1e6)) >= 2
    np.testing.assert_allclose(out["depth.ratio"], ref[keep, 3], rtol=RTOL)
    np.testing.assert_allclose(out["Bf"], ref[keep, 0], rtol=RTOL)


def test_segment_stats_large_genome_properties():
    """2e6 rows, 24 chromosomes, 1,200 segments: counts add up, a permutation-free split of
    the genome by chromosome gives the same rows, the BAF mode sits at the simulated value."""
    g = G.synth_seqz(n_chr=24, rows_per_chr=83333, segs_per_chr=50, seed=9)
    args = lambda d: (d["row_off"], d["seg_off"], d["position"], d["adjusted_ratio"],  # noqa: E731
                      d["depth_normal"], d["good_reads"], d["Af"], d["Bf"], d["is_het"],
                      d["breaks_vect"])
    full = G.segment_stats(*args(g))
    assert full.shape == (1200, 6) and full[:, 4].sum() <= g["position"].size
    assert (full[:, 1] <= full[:, 4]).all() and np.nanmax(full[:, 0]) <= 0.5
    c = 7
    r = slice(g["row_off"][c], g["row_off"][c + 1])
    one = {k: (v[r] if k not in ("row_off", "seg_off", "breaks_vect") else v) for k, v in g.items()}
    one["row_off"] = np.array([0, 83333])
    one["seg_off"] = np.array([0, 50])
    one["breaks_vect"] = g["breaks_vect"][100 * c:100 * (c + 1)]
    part = G.segment_stats(*args(one))
    np.testing.assert_array_equal(part, full[50 * c:50 * (c + 1)])
