"""The candidate rule of the pruned grid kernels, restated in numpy (CPU only).

For a segment and a tile of sorted ploidies the kernels bound q = wub (1 + x^2) of every
copy number from below, evaluate the one with the smallest bound, take Qmax >= its largest
q over the tile (one of the two tile ends, x being monotone) plus one step of 2^-20, and
skip every copy number whose bound is not below Qmax.  This test checks on random and on
awkward inputs that the skipped copy numbers never lower a minimum, i.e. that the pruned
minimum equals the full one for every (segment, ploidy), and that the exponent bound the
column-per-lane kernel uses to postpone renormalisation really bounds the products."""
import numpy as np
import pytest

from sequenza_b200.model import theoretical_baf, theoretical_depth_ratio
from sequenza_b200.synth import synth_segments


def column_terms(seg, cell, cn_max):
    """wub[c, s] = prior^(-1/3) min_B (1 + xb^2/5) as the pre-pass builds it (CN 2 has prior 2)."""
    n = seg["depth_ratio"].size
    hasb = ~np.isnan(seg["Bf"])
    kb = np.where(hasb, 1 / (seg["sd_Bf"] / np.sqrt(seg["weight_Bf"])) / np.sqrt(5), 0.0)
    eb = np.where(hasb, -seg["Bf"] * kb, 0.0)
    wub = np.empty((cn_max + 1, n))
    rows = np.array([c // 2 + 1 for c in range(cn_max + 1)])
    value = np.where(np.arange(cn_max + 1) == 2, 2.0, 1.0)
    prior = value / np.sum(rows * value)  # R/bayes.R:55-59: divided by the sum over type rows
    for c in range(cn_max + 1):
        mb = np.array([theoretical_baf(c, b, cell) for b in range(c // 2 + 1)])
        xb = mb[:, None] * kb[None, :] + eb[None, :]
        m0 = np.min(np.abs(xb), axis=0)
        wub[c] = (1 + m0 * m0) * prior[c] ** (-1 / 3)
    return wub


@pytest.mark.parametrize("cn_max", [7, 20])
@pytest.mark.parametrize("pt", [16, 32])
@pytest.mark.parametrize("scale", [0.05, 1.0, 20.0])
def test_pruned_minimum_equals_full_minimum(cn_max, pt, scale):
    rng = np.random.default_rng(cn_max * 100 + pt)
    seg = synth_segments(400, seed=11 + pt, cn_max=min(cn_max, 12), na_frac=0.05)
    seg["sd_ratio"] = seg["sd_ratio"] * scale
    seg["sd_Bf"] = seg["sd_Bf"] * scale
    kr = 1 / (seg["sd_ratio"] / np.sqrt(seg["weight_ratio"])) / np.sqrt(5)
    er = -np.log(seg["depth_ratio"]) * kr
    cn = np.arange(cn_max + 1)
    ploidy = np.sort(np.concatenate([np.arange(1, 7.001, 0.01), rng.uniform(0.5, 9, 7)]))
    grown_ok = True
    for cell in (0.05, 0.31, 0.7, 0.999, 1.0):
        wub = column_terms(seg, cell, cn_max)
        rw = np.sqrt(wub)
        with np.errstate(divide="ignore"):
            la = np.log(1.0 * ((1 - cell) + cn / 2.0 * cell))  # avg.depth.ratio = 1
        for p0 in rng.choice(ploidy.size - pt, size=6, replace=False):
            pj = ploidy[p0:p0 + pt]
            nld = -np.log(pj / 2.0 * cell + 1 - cell)
            nmid = 0.5 * (nld[0] + nld[-1])
            nld = nld - nmid
            hw = max(abs(nld[0]), abs(nld[-1]))
            lm = la + nmid
            # q[c, j, s] exactly as the kernel forms it: x = nld_j * krw + base_c
            krw = kr[None, :] * rw
            with np.errstate(invalid="ignore", over="ignore"):
                base = lm[:, None] * krw + er[None, :] * rw
                x = nld[None, :, None] * krw[:, None, :] + base[:, None, :]
                q = x * x + wub[:, None, :]
                full = np.fmin.reduce(np.where(np.isnan(q), np.inf, q), axis=0)  # NaN never wins
                # bounds
                t = np.abs(lm[:, None] * kr[None, :] + er[None, :]) - hw * np.abs(kr)[None, :]
                t = np.maximum(t, 0.0)
                lb = wub * (1 + t * t)
            lb = np.where(np.isnan(lb), np.inf, lb)
            first = np.argmin(lb, axis=0)
            s_idx = np.arange(kr.size)
            qf = q[first, :, s_idx].T  # [j, s]
            with np.errstate(invalid="ignore"):
                qend = np.fmax(qf[0], qf[-1])
            ok = np.isfinite(qend)
            qmax = np.where(ok, qend * (1 + 2.0 ** -20), np.inf)
            need = lb < qmax[None, :]
            need[first, s_idx] = True
            pruned = np.fmin.reduce(np.where(need[:, None, :] & ~np.isnan(q), q, np.inf), axis=0)
            # segments whose first candidate is NaN/Inf are flagged and repaired by the
            # faithful kernel; everywhere else the minima must agree exactly
            good = ok & np.isfinite(qf).all(axis=0)
            assert good.sum() > 0.8 * good.size
            np.testing.assert_array_equal(pruned[:, good], full[:, good])
            # the tile-end q of the first candidate bounds every factor (renormalisation rule)
            grown_ok &= bool((full[:, good] <= qend[good][None, :] * (1 + 2.0 ** -20)).all())
            assert (full[:, good] >= 1.0).all()  # products never underflow
    assert grown_ok


def test_depth_ratio_splits_into_copy_number_and_ploidy_terms():
    """log mr = log(adr A_c) - log D_j, the factorisation the column-per-lane kernel uses
    (R/model.R:1-7), to a few ulp of the table the other kernels build."""
    cn = np.arange(0, 21)[:, None, None]
    cell = np.array([0.05, 0.4, 0.999])[None, :, None]
    pl = np.arange(1, 7.001, 0.1)[None, None, :]
    adr = 0.93
    with np.errstate(divide="ignore"):
        direct = np.log(theoretical_depth_ratio(cn, cell, pl, avg_depth_ratio=adr))
        split = np.log(adr * ((1 - cell) + cn / 2.0 * cell)) - np.log(pl / 2.0 * cell + 1 - cell)
    fin = np.isfinite(direct)
    assert np.array_equal(fin, np.isfinite(split))
    assert np.max(np.abs(direct[fin] - split[fin])) < 1e-14
