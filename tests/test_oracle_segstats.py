"""The CPU oracle of segment.breaks' per-segment statistics (R/segments.R:113-198) and of
b_allele_freq (R/model.R:69-98) against an independent numpy restatement: density.default
as R does it (linear binning onto 1024 points, FFT correlation with dnorm, approx onto 512
points), bw.nrd0 from numpy's std / quantile, cut() as interval tests.  CPU only.  (The
reference's own tests never call these functions: parity stays unpinned against R.)"""
import numpy as np
import pytest


def np_density_mode(Af, Bf, good):
    """b_allele_freq(...)[2] with numpy: returns (mode, density values on the 512 grid)."""
    m = len(Bf)
    if m == 0:
        return np.nan, None
    if m == 1:
        return Bf[0], None
    x = np.concatenate([Bf, Af])
    w = np.concatenate([good, good]) / (2 * good.sum())
    n = 512
    hi = np.std(x, ddof=1)
    q25, q75 = np.quantile(x, [0.25, 0.75])  # type 7
    lo = min(hi, (q75 - q25) / 1.34) or hi or abs(x[0]) or 1.0
    bw = 0.9 * lo * x.size ** -0.2
    frm, to = x.min() - 3 * bw, x.max() + 3 * bw
    a, b = frm - 4 * bw, to + 4 * bw
    y = np.zeros(2 * n)
    xpos = (x - a) / ((b - a) / (n - 1))
    ix = np.floor(xpos).astype(int)
    fx = xpos - ix
    ok = (ix >= 0) & (ix <= n - 2)
    np.add.at(y, ix[ok], w[ok] * (1 - fx[ok]))
    np.add.at(y, ix[ok] + 1, w[ok] * fx[ok])
    np.add.at(y, np.zeros((ix == -1).sum(), dtype=int), (w * fx)[ix == -1])
    np.add.at(y, ix[ix == n - 1], (w * (1 - fx))[ix == n - 1])
    k = np.linspace(0, 2 * (b - a), 2 * n)
    k[n + 1:] = -k[n - 1:0:-1]
    k = np.exp(-0.5 * (k / bw) ** 2) / (bw * np.sqrt(2 * np.pi))
    conv = np.fft.ifft(np.fft.fft(y) * np.conj(np.fft.fft(k))).real[:n]
    conv = np.maximum(0, conv)
    gx = np.linspace(frm, to, n)
    gy = np.interp(gx, np.linspace(a, b, n), conv)
    # local maxima, a run of equal values at the top counting as one (at its first point):
    # see the note in oracle/sqz_oracle_segstats.c
    sg = np.sign(np.diff(gy))
    pk, cand = [], -1
    for j, v in enumerate(sg):
        if v > 0:
            cand = j + 1
        elif v < 0:
            if cand >= 0:
                pk.append(cand)
            cand = -1
    pk = np.array(pk, dtype=int)
    if pk.size == 0:
        pk = np.flatnonzero(np.diff(sg) == -1) + 1
    best = pk[np.argmax(gy[pk])]
    v = gx[best]
    return (1 - v if v > 0.5 else v), gy


def synth_het(rng, m, baf, depth=60):
    g = np.round(rng.uniform(0.5 * depth, 1.5 * depth, m))
    b = rng.binomial(g.astype(int), baf) / g
    b = np.minimum(b, 1 - b)
    return 1 - b, b, g


@pytest.mark.parametrize("m,baf", [(2, 0.3), (3, 0.5), (40, 0.33), (500, 0.5), (5000, 0.12),
                                   (20000, 0.41)])
def test_density_mode_matches_numpy_fft(oracle, m, baf):
    rng = np.random.default_rng(m)
    Af, Bf, g = synth_het(rng, m, baf)
    got, dens = oracle.b_allele_freq_mode(Af, Bf, g, return_density=True)
    want, gy = np_density_mode(Af, Bf, g)
    # the FFT leaves ~1e-17 noise where the density is zero; compare where it is not
    big = gy > 1e-9 * gy.max()
    np.testing.assert_allclose(dens[big], gy[big], rtol=1e-9)
    assert got == pytest.approx(want, rel=1e-12)
    if m >= 40:
        assert abs(got - min(baf, 1 - baf)) < 0.06


def test_density_mode_edge_cases(oracle):
    assert np.isnan(oracle.b_allele_freq_mode([], [], []))
    assert oracle.b_allele_freq_mode([0.7], [0.3], [50.0]) == 0.3       # length(Bf) == 1
    # all values equal: sd = IQR = 0 -> bw.nrd0 falls back to |x[1]|
    v, _ = oracle.b_allele_freq_mode([0.5] * 6, [0.5] * 6, [30.0] * 6, return_density=True)
    w, _ = np_density_mode(np.full(6, 0.5), np.full(6, 0.5), np.full(6, 30.0))
    assert v == pytest.approx(w, rel=1e-12)


def np_segment_stats(pos, ratio, dn, gr, Af, Bf, het, brk, min_reads=1, weighted=True):
    out = []
    for s in range(len(brk) // 2):
        b0, b1 = brk[2 * s], brk[2 * s + 1]
        inr = (pos > b0) & (pos <= b1)                      # cut(): (b0, b1]
        hb = inr & het & (gr >= min_reads)
        if weighted:
            wr = np.sqrt(dn[inr])
            rw = ratio[inr] * wr
            dr = np.nansum(rw) / np.nansum(wr)
            v = rw / wr
            wb = np.sqrt(gr[hb])
            vb = (Bf[hb] * wb) / wb
        else:
            v = ratio[inr]
            dr = np.nanmean(v) if np.isfinite(v).any() else np.nan
            vb = Bf[hb]
        v = v[~np.isnan(v)]
        sd = lambda a: np.std(a, ddof=1) if a.size > 1 else np.nan  # noqa: E731
        mode, _ = np_density_mode(Af[hb], Bf[hb], gr[hb])
        out.append([mode, hb.sum(), sd(vb), dr, inr.sum(), sd(v)])
    return np.array(out)


def make_genome(rng, n_chr=3, rows=4000, segs=7):
    from sequenza_b200.segments import synth_seqz
    return synth_seqz(n_chr=n_chr, rows_per_chr=rows, segs_per_chr=segs, seed=int(rng.integers(1e9)))


@pytest.mark.parametrize("weighted", [True, False])
def test_segment_stats_match_numpy(oracle, weighted):
    rng = np.random.default_rng(7)
    g = make_genome(rng)
    got = oracle.segment_stats(g["row_off"], g["seg_off"], g["position"], g["adjusted_ratio"],
                               g["depth_normal"], g["good_reads"], g["Af"], g["Bf"], g["is_het"],
                               g["breaks_vect"], 1, weighted)
    want = []
    for c in range(len(g["row_off"]) - 1):
        r = slice(g["row_off"][c], g["row_off"][c + 1])
        b = g["breaks_vect"][2 * g["seg_off"][c]:2 * g["seg_off"][c + 1]]
        want.append(np_segment_stats(g["position"][r], g["adjusted_ratio"][r],
                                     g["depth_normal"][r], g["good_reads"][r], g["Af"][r],
                                     g["Bf"][r], g["is_het"][r].astype(bool), b, 1, weighted))
    want = np.vstack(want)
    assert got.shape == want.shape == (g["seg_off"][-1], 6)
    np.testing.assert_array_equal(got[:, [1, 4]], want[:, [1, 4]])          # N.BAF, N.ratio
    for col in (0, 2, 3, 5):
        a, b = got[:, col], want[:, col]
        assert np.array_equal(np.isnan(a), np.isnan(b)), col
        ok = ~np.isnan(a)
        np.testing.assert_allclose(a[ok], b[ok], rtol=1e-10, err_msg="column %d" % col)
    # a position equal to a segment's start is NOT in it, one equal to its end is (cut)
    assert (got[:, 4] > 0).any()
