"""ctypes front-end of the CPU oracle (oracle/sqz_oracle.c) plus the few R-level steps
that stay in R in the reference (tapply/normalise, get.ci).

TEST INFRASTRUCTURE ONLY -- see the header of sqz_oracle.c.  The product package
``sequenza_b200`` never imports this module.

Function names and arguments mirror the reference (R/bayes.R, R/model.R,
R/solutions.R) with dots turned into underscores.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libsqz_oracle.so")
_lib = None

NA_REAL_BITS = 0x7FF00000000007A2


def build(force: bool = False) -> str:
    """Compile the oracle with the committed Makefile (gcc, no R needed)."""
    srcs = [os.path.join(_HERE, f) for f in ("sqz_oracle.c", "sqz_oracle_segstats.c")]
    if force or not os.path.exists(_SO) or \
            os.path.getmtime(_SO) < max(os.path.getmtime(f) for f in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B", "libsqz_oracle.so"])
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        d, i, i64, p = C.c_double, C.c_int, C.c_int64, C.c_void_p
        _lib.sqzo_na_real.restype = d
        for name, args in {
            "sqzo_stirlerr": [d],
            "sqzo_bd0": [d, d],
            "sqzo_round0": [d],
            "sqzo_theoretical_depth_ratio": [d] * 6,
            "sqzo_theoretical_baf": [d] * 4,
            "sqzo_theoretical_mufreq": [d] * 4,
        }.items():
            getattr(_lib, name).restype = d
            getattr(_lib, name).argtypes = args
        _lib.sqzo_dt.restype = d
        _lib.sqzo_dt.argtypes = [d, d, i]
        _lib.sqzo_dbinom.restype = d
        _lib.sqzo_dbinom.argtypes = [d, d, d, i]
        _lib.sqzo_baf_types_matrix.restype = i
        _lib.sqzo_baf_types_matrix.argtypes = [i, i, i, p]
        _lib.sqzo_mufreq_types_matrix.restype = i
        _lib.sqzo_mufreq_types_matrix.argtypes = [i, i, i, p]
        _lib.sqzo_baf_bayes.restype = i
        _lib.sqzo_baf_bayes.argtypes = [i64] + [p] * 6 + [d, d, d, i, i, i, i, p, p, i, p, p]
        _lib.sqzo_baf_model_fit_sums.restype = i
        _lib.sqzo_baf_model_fit_sums.argtypes = (
            [i64] + [p] * 6 + [p, i64, p, i64, d, i, i, i, i, p, p, i, i64, i64, i, p]
        )
        _lib.sqzo_mufreq_bayes.restype = i
        _lib.sqzo_mufreq_bayes.argtypes = [i64] + [p] * 4 + [d, d, d, i, i, i, i, p, p, p]
        _lib.sqzo_mufreq_model_fit_sums.restype = i
        _lib.sqzo_mufreq_model_fit_sums.argtypes = (
            [i64] + [p] * 4 + [p, i64, p, i64, d, i, i, i, i, p, p, i64, i64, i, p]
        )
        _lib.sqzo_normalise_lpp.restype = None
        _lib.sqzo_normalise_lpp.argtypes = [i64, p, p]
        _lib.sqzo_max_threads.restype = i
        _lib.sqzo_b_allele_freq_mode.restype = d
        _lib.sqzo_b_allele_freq_mode.argtypes = [p, p, p, i64, p]
        _lib.sqzo_segment_stats.restype = i
        _lib.sqzo_segment_stats.argtypes = [i64, p, p] + [p] * 8 + [d, i, p]
    return _lib


def _vec(x, n=None):
    a = np.ascontiguousarray(np.asarray(x, dtype=np.float64).reshape(-1))
    if n is not None and a.size != n:
        if a.size == 1:
            a = np.full(n, a[0], dtype=np.float64)  # R's scalar recycling
        else:
            raise ValueError("length mismatch: %d vs %d" % (a.size, n))
    return a


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def _priors(priors_table, CNt_min, CNt_max):
    """priors.table as (CN, value) pairs; default R/bayes.R:40-41 (all ones)."""
    if priors_table is None:
        cn = np.arange(CNt_min, CNt_max + 1, dtype=np.float64)
        return cn, np.ones_like(cn)
    if isinstance(priors_table, dict):
        cn, val = priors_table["CN"], priors_table["value"]
    else:
        cn, val = priors_table
    return _vec(cn), _vec(val)


def is_na(x):
    return np.isnan(x)


# ---- nmath / model scalars -------------------------------------------------
def dt(x, df, log=False):
    return lib().sqzo_dt(float(x), float(df), int(log))


def dbinom(x, size, prob, log=False):
    return lib().sqzo_dbinom(float(x), float(size), float(prob), int(log))


def theoretical_depth_ratio(CNt, cellularity, ploidy, CNn=2, normal_ploidy=2,
                            avg_depth_ratio=1):
    return lib().sqzo_theoretical_depth_ratio(CNt, cellularity, ploidy, CNn,
                                              normal_ploidy, avg_depth_ratio)


def theoretical_baf(CNt, B, cellularity, CNn=2):
    return lib().sqzo_theoretical_baf(CNt, B, cellularity, CNn)


def theoretical_mufreq(CNt, Mt, cellularity, CNn=2):
    return lib().sqzo_theoretical_mufreq(CNt, Mt, cellularity, CNn)


def baf_types_matrix(CNt_min, CNt_max, CNn=2):
    T = lib().sqzo_baf_types_matrix(CNt_min, CNt_max, CNn, None)
    out = np.empty((3, T))
    lib().sqzo_baf_types_matrix(CNt_min, CNt_max, CNn, _ptr(out))
    return out.T.copy()  # columns CNn, CNt, B


def mufreq_types_matrix(CNt_min, CNt_max, CNn=2):
    T = lib().sqzo_mufreq_types_matrix(CNt_min, CNt_max, CNn, None)
    out = np.empty((3, T))
    lib().sqzo_mufreq_types_matrix(CNt_min, CNt_max, CNn, _ptr(out))
    return out.T.copy()  # columns CNn, CNt, Mt


# ---- R/bayes.R ---------------------------------------------------------------
def baf_bayes(Bf, depth_ratio, cellularity, ploidy, avg_depth_ratio, sd_Bf=0.1,
              sd_ratio=0.5, weight_Bf=1, weight_ratio=1, CNt_min=0, CNt_max=7, CNn=2,
              priors_table=None, ratio_priority=False, return_margin=False):
    """R/bayes.R:38-101.  Returns an N x 4 array, columns CNt, A, B, LPP."""
    r = _vec(depth_ratio)
    n = r.size
    args = [_vec(v, n) for v in (Bf, r, sd_Bf, sd_ratio, weight_Bf, weight_ratio)]
    pc, pv = _priors(priors_table, CNt_min, CNt_max)
    out = np.empty((4, n))
    margin = np.empty(n)
    rc = lib().sqzo_baf_bayes(n, *[_ptr(a) for a in args], cellularity, ploidy,
                              avg_depth_ratio, CNt_min, CNt_max, CNn, pc.size, _ptr(pc),
                              _ptr(pv), int(bool(ratio_priority)), _ptr(out),
                              _ptr(margin))
    if rc:
        raise MemoryError("oracle baf_bayes")
    res = out.T.copy()
    return (res, margin) if return_margin else res


def baf_grid_sums(Bf, depth_ratio, cellularity, ploidy, avg_depth_ratio, sd_Bf=0.1,
                  sd_ratio=0.5, weight_Bf=1, weight_ratio=1, CNt_min=0, CNt_max=7, CNn=2,
                  priors_table=None, ratio_priority=False, g0=0, g1=None, nthreads=0):
    """R/bayes.R:106-116: result$LPP in expand.grid order (ploidy fastest)."""
    r = _vec(depth_ratio)
    n = r.size
    args = [_vec(v, n) for v in (Bf, r, sd_Bf, sd_ratio, weight_Bf, weight_ratio)]
    pl, ce = _vec(ploidy), _vec(cellularity)
    pc, pv = _priors(priors_table, CNt_min, CNt_max)
    G = pl.size * ce.size
    g1 = G if g1 is None else g1
    out = np.empty(g1 - g0)
    rc = lib().sqzo_baf_model_fit_sums(n, *[_ptr(a) for a in args], _ptr(pl), pl.size,
                                       _ptr(ce), ce.size, avg_depth_ratio, CNt_min,
                                       CNt_max, CNn, pc.size, _ptr(pc), _ptr(pv),
                                       int(bool(ratio_priority)), g0, g1, nthreads,
                                       _ptr(out))
    if rc:
        raise MemoryError("oracle baf_model_fit_sums")
    return out


def _r_num_label(x):
    """as.numeric(as.character(x)): R prints 15 significant digits (R/bayes.R:117-119)."""
    return float("%.15g" % x)


def finish_model_fit(lpp_sum, ploidy, cellularity):
    """R/bayes.R:116-123: tapply(mean) into [ploidy x cellularity], axis values through
    their character dimnames, LogSumLik over the un-aggregated vector, exp(z - LSL)."""
    pl, ce = _vec(ploidy), _vec(cellularity)
    lpp_sum = _vec(lpp_sum, pl.size * ce.size)
    plab = [_r_num_label(v) for v in pl]
    clab = [_r_num_label(v) for v in ce]
    x = sorted(set(plab))
    y = sorted(set(clab))
    pi = np.array([x.index(v) for v in plab])
    ci = np.array([y.index(v) for v in clab])
    acc = np.zeros((len(x), len(y)), dtype=np.longdouble)
    cnt = np.zeros((len(x), len(y)))
    grid = lpp_sum.reshape(ce.size, pl.size)  # ploidy fastest
    for jc in range(ce.size):
        for jp in range(pl.size):
            acc[pi[jp], ci[jc]] += grid[jc, jp]
            cnt[pi[jp], ci[jc]] += 1
    z = (acc / cnt).astype(np.float64)
    # LogSumLik from the un-aggregated vector; then exp(z - LogSumLik) on the matrix
    finite = lpp_sum[~np.isnan(lpp_sum)]
    mx = finite.max() if finite.size else np.nan
    lsl = float(np.log(np.float64(np.sum(np.exp(lpp_sum - mx).astype(np.longdouble))))
                + mx)
    return {"ploidy": np.array(x), "cellularity": np.array(y), "lpp": np.exp(z - lsl)}


def baf_model_fit(cellularity=None, ploidy=None, mc_cores=2, **kw):
    """R/bayes.R:103-124."""
    if cellularity is None:
        cellularity = np.arange(30, 101) / 100.0
    if ploidy is None:
        ploidy = np.arange(10, 71) / 10.0
    sums = baf_grid_sums(cellularity=cellularity, ploidy=ploidy, nthreads=mc_cores, **kw)
    return finish_model_fit(sums, ploidy, cellularity)


def mufreq_bayes(mufreq, depth_ratio, cellularity, ploidy, avg_depth_ratio,
                 weight_mufreq=100, weight_ratio=100, CNt_min=1, CNt_max=7, CNn=2,
                 priors_table=None):
    """R/bayes.R:126-173.  Returns N x 4, columns CNn, CNt, Mt, LPP."""
    f = _vec(mufreq)
    n = f.size
    args = [_vec(v, n) for v in (f, depth_ratio, weight_mufreq, weight_ratio)]
    pc, pv = _priors(priors_table, CNt_min, CNt_max)
    out = np.empty((4, n))
    rc = lib().sqzo_mufreq_bayes(n, *[_ptr(a) for a in args], cellularity, ploidy,
                                 avg_depth_ratio, CNt_min, CNt_max, CNn, pc.size,
                                 _ptr(pc), _ptr(pv), _ptr(out))
    if rc:
        raise MemoryError("oracle mufreq_bayes")
    return out.T.copy()


def mufreq_grid_sums(mufreq, depth_ratio, cellularity, ploidy, avg_depth_ratio,
                     weight_mufreq=100, weight_ratio=100, CNt_min=1, CNt_max=7, CNn=2,
                     priors_table=None, g0=0, g1=None, nthreads=0):
    """R/bayes.R:177-186."""
    f = _vec(mufreq)
    n = f.size
    args = [_vec(v, n) for v in (f, depth_ratio, weight_mufreq, weight_ratio)]
    pl, ce = _vec(ploidy), _vec(cellularity)
    pc, pv = _priors(priors_table, CNt_min, CNt_max)
    G = pl.size * ce.size
    g1 = G if g1 is None else g1
    out = np.empty(g1 - g0)
    rc = lib().sqzo_mufreq_model_fit_sums(n, *[_ptr(a) for a in args], _ptr(pl), pl.size,
                                          _ptr(ce), ce.size, avg_depth_ratio, CNt_min,
                                          CNt_max, CNn, pc.size, _ptr(pc), _ptr(pv), g0,
                                          g1, nthreads, _ptr(out))
    if rc:
        raise MemoryError("oracle mufreq_model_fit_sums")
    return out


def mufreq_model_fit(cellularity=None, ploidy=None, mc_cores=2, **kw):
    """R/bayes.R:175-194."""
    if cellularity is None:
        cellularity = np.arange(30, 101) / 100.0
    if ploidy is None:
        ploidy = np.arange(10, 71) / 10.0
    sums = mufreq_grid_sums(cellularity=cellularity, ploidy=ploidy, nthreads=mc_cores,
                            **kw)
    return finish_model_fit(sums, ploidy, cellularity)


# ---- R/solutions.R -------------------------------------------------------------
def get_ci(cp_table, level=0.95):
    """R/solutions.R:1-26, written as plain loops (small inputs only)."""
    lpp = np.asarray(cp_table["lpp"], dtype=np.float64)
    ploidy = np.asarray(cp_table["ploidy"], dtype=np.float64)
    cell = np.asarray(cp_table["cellularity"], dtype=np.float64)
    flat = [v for v in lpp.reshape(-1) if not np.isnan(v)]  # sort() drops NA
    flat.sort(reverse=True)
    run = np.longdouble(0)
    n_below = 0
    for v in flat:  # cumsum: long double accumulator, stored as double
        run += v
        if float(run) < level:
            n_below += 1
    thresh = flat[min(n_below, len(flat) - 1)]
    vx = [max(lpp[i, :]) for i in range(lpp.shape[0])]  # apply(lpp, 1, max)
    vy = [max(lpp[:, j]) for j in range(lpp.shape[1])]  # apply(lpp, 2, max)
    px = [ploidy[i] for i in range(len(vx)) if vx[i] >= thresh]
    cy = [cell[j] for j in range(len(vy)) if vy[j] >= thresh]

    def first_max(v):
        k, best = -1, None
        for i, a in enumerate(v):
            if not np.isnan(a) and (k < 0 or a > best):
                k, best = i, a
        return k

    return {
        "values.ploidy": {"x": ploidy.copy(), "y": np.array(vx) / float(np.sum(vx))},
        "confint.ploidy": (min(px), max(px)),
        "max.ploidy": ploidy[first_max(vx)],
        "values.cellularity": {"x": np.array(vy) / float(np.sum(vy)), "y": cell.copy()},
        "confint.cellularity": (min(cy), max(cy)),
        "max.cellularity": cell[first_max(vy)],
    }


# ---- R/segments.R, R/model.R:69-98 -----------------------------------------------
def b_allele_freq_mode(Af, Bf, good_reads, return_density=False):
    """b_allele_freq(Af, Bf, good.reads)[2] (R/model.R:69-98): the mode of the weighted
    kernel density of c(Bf, Af), folded to <= 0.5."""
    a, b, g = _vec(Af), _vec(Bf), _vec(good_reads)
    dens = np.empty(512)
    v = lib().sqzo_b_allele_freq_mode(_ptr(a), _ptr(b), _ptr(g), b.size, _ptr(dens))
    return (v, dens) if return_density else v


SEGMENT_COLUMNS = ("Bf", "N.BAF", "sd.BAF", "depth.ratio", "N.ratio", "sd.ratio")


def segment_stats(row_off, seg_off, position, adjusted_ratio, depth_normal, good_reads, Af,
                  Bf, is_het, breaks_vect, min_reads_baf=1, weighted_mean=True):
    """The per-segment statistics of segment.breaks (R/segments.R:113-198) for a genome laid
    out as concatenated chromosomes; returns an n_seg x 6 array, columns SEGMENT_COLUMNS."""
    ro = np.ascontiguousarray(row_off, dtype=np.int64)
    so = np.ascontiguousarray(seg_off, dtype=np.int64)
    cols = [_vec(v) for v in (position, adjusted_ratio, depth_normal, good_reads, Af, Bf)]
    het = np.ascontiguousarray(is_het, dtype=np.uint8)
    brk = _vec(breaks_vect)
    S = int(so[-1])
    out = np.empty((6, S))
    rc = lib().sqzo_segment_stats(ro.size - 1, _ptr(ro), _ptr(so), *[_ptr(c) for c in cols],
                                  _ptr(het), _ptr(brk), float(min_reads_baf),
                                  int(bool(weighted_mean)), _ptr(out))
    if rc:
        raise MemoryError("oracle segment_stats")
    return out.T.copy()
