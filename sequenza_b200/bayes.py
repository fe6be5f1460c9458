"""Drop-in mirrors of R/bayes.R's public functions over the CUDA library.

Argument names follow the reference with '.' turned into '_'.  Every function here
runs on the GPU through the C ABI (include/sequenza_b200.h); nothing in this module
computes likelihoods on the host.
"""
from __future__ import annotations

import numpy as np

from . import _lib
from ._lib import SQZ_MODE_FAITHFUL, SQZ_MODE_FAST  # noqa: F401  (re-exported)


def r_seq(frm, to, by):
    """R's seq(from, to, by) for by > 0 (seq.default: from + (0:n)*by, 1e-10 fuzz)."""
    n = int(np.floor((to - frm) / by + 1e-10))
    return np.minimum(frm + np.arange(n + 1) * by, to)


def _priors(priors_table, CNt_min, CNt_max):
    """priors.table -> (CN, value) float64 vectors.  Default: R/bayes.R:40-41, :128-129."""
    if priors_table is None:
        cn = np.arange(CNt_min, CNt_max + 1, dtype=np.float64)
        return cn, np.ones_like(cn)
    if hasattr(priors_table, "columns"):  # pandas.DataFrame
        cn, val = priors_table["CN"].to_numpy(), priors_table["value"].to_numpy()
    elif isinstance(priors_table, dict):
        cn, val = priors_table["CN"], priors_table["value"]
    else:
        cn, val = priors_table
    cn, val = _lib.vec(cn), _lib.vec(val)
    if cn.size != val.size:
        raise ValueError("priors.table: CN and value differ in length")
    return cn, val


def _r_label(x):
    """as.numeric(as.character(x)): 15 significant digits (R/bayes.R:118-119)."""
    return float("%.15g" % x)


class CpTable(dict):
    """list(ploidy, cellularity, lpp) as a dict.  When the device produced it
    (sqz_fit_finish) it also carries, outside the mapping, the reductions get.ci needs, so
    that get_ci() does not have to sort the matrix again."""
    ci_cache = None


def finish_model_fit_device(lpp_sum, x, y):
    """R/bayes.R:120-122 and the get.ci reductions on the GPU (sqz_fit_finish) for sorted,
    unique axis values x (ploidy) and y (cellularity); lpp_sum may hold several samples."""
    nP, nC = x.size, y.size
    lpp_sum = np.ascontiguousarray(lpp_sum, dtype=np.float64).reshape(-1, nP * nC)
    S = lpp_sum.shape[0]
    lpp = np.empty((S, nP * nC))
    rmax, rab = np.empty((S, nP)), np.empty((S, nP))
    cmax, cab = np.empty((S, nC)), np.empty((S, nC))
    lsl = np.empty(S)
    _lib.check(_lib.lib().sqz_fit_finish(
        _lib.ptr(lpp_sum), S, nP, nC, _lib.ptr(lpp), _lib.ptr(rmax), _lib.ptr(cmax),
        _lib.ptr(rab), _lib.ptr(cab), _lib.ptr(lsl)))
    out = []
    for k in range(S):
        t = CpTable(ploidy=x.copy(), cellularity=y.copy(),
                    lpp=lpp[k].reshape(nC, nP).T)  # R matrix [ploidy x cellularity]
        t.ci_cache = {"lpp_id": id(t["lpp"]), "row_max": rmax[k], "col_max": cmax[k],
                      "row_above": rab[k], "col_above": cab[k], "log_sum_lik": lsl[k]}
        out.append(t)
    return out


def finish_model_fit(lpp_sum, ploidy, cellularity, device=None):
    """R/bayes.R:116-123 on the length-G vector the GPU returned: tapply(mean) into a
    [ploidy x cellularity] matrix whose axes went through character dimnames, then
    LogSumLik (max with na.rm, sum without) and exp(z - LogSumLik).
    device=None: use the GPU (sqz_fit_finish) when the axes are sorted and unique, there is
    no NaN and a device is present; False: numpy (the reference's steps, one by one)."""
    pl, ce = _lib.vec(ploidy), _lib.vec(cellularity)
    lpp_sum = _lib.vec(lpp_sum, pl.size * ce.size, "lpp_sum")
    x, pi = np.unique(np.array([_r_label(v) for v in pl]), return_inverse=True)
    y, ci = np.unique(np.array([_r_label(v) for v in ce]), return_inverse=True)
    if device is None or device:
        sorted_unique = (x.size == pl.size and y.size == ce.size and
                         np.array_equal(pi, np.arange(pl.size)) and
                         np.array_equal(ci, np.arange(ce.size)))
        usable = (sorted_unique and pl.size + ce.size <= 4096 and pl.size * ce.size > 0 and
                  not np.isnan(lpp_sum).any())
        if usable and (device or _lib.lib().sqz_device_count() > 0):
            return finish_model_fit_device(lpp_sum, x, y)[0]
        if device:
            raise ValueError("device finish needs sorted unique axes, no NaN, <= 4096 values")
    grid = lpp_sum.reshape(ce.size, pl.size)  # expand.grid: ploidy varies fastest
    if x.size == pl.size and y.size == ce.size:
        z = np.empty((x.size, y.size))
        z[np.ix_(pi, ci)] = grid.T
    else:  # duplicated axis values: tapply averages them
        acc = np.zeros((x.size, y.size), dtype=np.longdouble)
        cnt = np.zeros((x.size, y.size))
        np.add.at(acc, (pi[None, :].repeat(ce.size, 0), ci[:, None].repeat(pl.size, 1)),
                  grid.astype(np.longdouble))
        np.add.at(cnt, (pi[None, :].repeat(ce.size, 0), ci[:, None].repeat(pl.size, 1)), 1)
        z = (acc / cnt).astype(np.float64)
    ok = ~np.isnan(lpp_sum)
    max_lik = lpp_sum[ok].max() if ok.any() else np.nan
    with np.errstate(invalid="ignore", over="ignore"):
        terms = np.exp(lpp_sum - max_lik).astype(np.longdouble)
        total = np.cumsum(terms)[-1] if terms.size else np.longdouble(0)  # sequential, as R sum()
        log_sum_lik = np.log(np.float64(total)) + max_lik
        znorm = np.exp(z - log_sum_lik)
    return {"ploidy": x, "cellularity": y, "lpp": znorm}


def baf_grid_sums(Bf, depth_ratio, cellularity, ploidy, avg_depth_ratio, sd_Bf=0.1,
                  sd_ratio=0.5, weight_Bf=1, weight_ratio=1, CNt_min=0, CNt_max=7, CNn=2,
                  priors_table=None, ratio_priority=False, n_gpus=1, mode=SQZ_MODE_FAST):
    """result$LPP of R/bayes.R:116 (ploidy fastest), from sqz_baf_grid."""
    r = _lib.vec(depth_ratio)
    n = r.size
    v = [_lib.vec(a, n, nm) for a, nm in ((Bf, "Bf"), (r, "depth.ratio"), (sd_Bf, "sd.Bf"),
                                          (sd_ratio, "sd.ratio"), (weight_Bf, "weight.Bf"),
                                          (weight_ratio, "weight.ratio"))]
    pl, ce = _lib.vec(ploidy), _lib.vec(cellularity)
    pc, pv = _priors(priors_table, CNt_min, CNt_max)
    out = np.empty(pl.size * ce.size)
    _lib.check(_lib.lib().sqz_baf_grid(
        *[_lib.ptr(a) for a in v], n, _lib.ptr(pl), pl.size, _lib.ptr(ce), ce.size,
        float(avg_depth_ratio), int(CNt_min), int(CNt_max), int(CNn), _lib.ptr(pc),
        _lib.ptr(pv), pc.size, int(bool(ratio_priority)), int(n_gpus), int(mode),
        _lib.ptr(out)))
    return out


def baf_model_fit(cellularity=None, ploidy=None, mc_cores=2, n_gpus=None, **kwargs):
    """R/bayes.R:103-124.  `mc_cores` is accepted for signature compatibility and
    ignored (there is no fork pool); `n_gpus` (default: all visible devices) is how many
    GPUs of this process share the cellularity columns."""
    cellularity = r_seq(0.3, 1, 0.01) if cellularity is None else cellularity
    ploidy = r_seq(1, 7, 0.1) if ploidy is None else ploidy
    if n_gpus is None:
        n_gpus = max(1, _lib.lib().sqz_device_count())
    sums = baf_grid_sums(cellularity=cellularity, ploidy=ploidy, n_gpus=n_gpus, **kwargs)
    return finish_model_fit(sums, ploidy, cellularity)


def baf_model_fit_batch(samples, cellularity=None, ploidy=None, n_gpus=None, CNt_min=0,
                        CNt_max=7, CNn=2, priors_table=None, ratio_priority=False,
                        mode=SQZ_MODE_FAST):
    """Cohort form (BASELINE config 5): `samples` is a list of dicts with the baf.model.fit
    per-sample arguments (Bf, depth_ratio, sd_Bf, sd_ratio, weight_Bf, weight_ratio,
    avg_depth_ratio); all samples share the grid.  A cohort that is already laid out as
    concatenated vectors can be passed as one dict with the same keys plus `seg_offsets`
    (S + 1 values) and a per-sample `avg_depth_ratio` vector.  One launch per device;
    returns one list(ploidy, cellularity, lpp) per sample."""
    cellularity = r_seq(0.3, 1, 0.01) if cellularity is None else cellularity
    ploidy = r_seq(1, 7, 0.1) if ploidy is None else ploidy
    pl, ce = _lib.vec(ploidy), _lib.vec(cellularity)
    keys = ("Bf", "depth_ratio", "sd_Bf", "sd_ratio", "weight_Bf", "weight_ratio")
    defaults = {"sd_Bf": 0.1, "sd_ratio": 0.5, "weight_Bf": 1, "weight_ratio": 1}
    if isinstance(samples, dict):  # pre-concatenated cohort
        offs = np.ascontiguousarray(samples["seg_offsets"], dtype=np.int64)
        n_tot = int(offs[-1]) if offs.size else 0
        cat = [_lib.vec(samples.get(k, defaults.get(k)), n_tot, k) for k in keys]
        adr = _lib.vec(samples["avg_depth_ratio"], max(offs.size - 1, 0), "avg.depth.ratio")
        n_samples = max(offs.size - 1, 0)
    else:
        cols = {k: [] for k in keys}
        off = [0]
        for smp in samples:
            n = _lib.vec(smp["depth_ratio"]).size
            for k in keys:
                cols[k].append(_lib.vec(smp.get(k, defaults.get(k)), n, k))
            off.append(off[-1] + n)
        cat = [np.concatenate(cols[k]) if cols[k] else np.empty(0) for k in keys]
        offs = np.asarray(off, dtype=np.int64)
        adr = _lib.vec([s["avg_depth_ratio"] for s in samples])
        n_samples = len(samples)
    pc, pv = _priors(priors_table, CNt_min, CNt_max)
    S, G = n_samples, pl.size * ce.size
    out = np.empty(S * G)
    if n_gpus is None:
        n_gpus = max(1, _lib.lib().sqz_device_count())
    _lib.check(_lib.lib().sqz_baf_grid_batch(
        S, offs.ctypes.data, *[_lib.ptr(a) for a in cat], _lib.ptr(pl), pl.size,
        _lib.ptr(ce), ce.size, _lib.ptr(adr), int(CNt_min), int(CNt_max), int(CNn),
        _lib.ptr(pc), _lib.ptr(pv), pc.size, int(bool(ratio_priority)), int(n_gpus),
        int(mode), _lib.ptr(out)))
    x = np.array([_r_label(v) for v in pl])
    y = np.array([_r_label(v) for v in ce])
    if (S > 0 and np.all(np.diff(x) > 0) and np.all(np.diff(y) > 0) and
            pl.size + ce.size <= 4096 and not np.isnan(out).any()):
        return finish_model_fit_device(out, x, y)  # one launch for the whole cohort
    return [finish_model_fit(out[k * G:(k + 1) * G], pl, ce) for k in range(S)]


def baf_bayes(Bf, depth_ratio, cellularity, ploidy, avg_depth_ratio, sd_Bf=0.1,
              sd_ratio=0.5, weight_Bf=1, weight_ratio=1, CNt_min=0, CNt_max=7, CNn=2,
              priors_table=None, ratio_priority=False):
    """R/bayes.R:38-101 -> N x 4 array with columns CNt, A, B, LPP (R/bayes.R:99)."""
    r = _lib.vec(depth_ratio)
    n = r.size
    v = [_lib.vec(a, n, nm) for a, nm in ((Bf, "Bf"), (r, "depth.ratio"), (sd_Bf, "sd.Bf"),
                                          (sd_ratio, "sd.ratio"), (weight_Bf, "weight.Bf"),
                                          (weight_ratio, "weight.ratio"))]
    pc, pv = _priors(priors_table, CNt_min, CNt_max)
    out = np.empty((4, n))
    _lib.check(_lib.lib().sqz_baf_point(
        *[_lib.ptr(a) for a in v], n, float(cellularity), float(ploidy),
        float(avg_depth_ratio), int(CNt_min), int(CNt_max), int(CNn), _lib.ptr(pc),
        _lib.ptr(pv), pc.size, int(bool(ratio_priority)), _lib.ptr(out)))
    return out.T.copy()


BAF_BAYES_COLUMNS = ("CNt", "A", "B", "LPP")
MUFREQ_BAYES_COLUMNS = ("CNn", "CNt", "Mt", "LPP")


def mufreq_bayes(mufreq, depth_ratio, cellularity, ploidy, avg_depth_ratio,
                 weight_mufreq=100, weight_ratio=100, CNt_min=1, CNt_max=7, CNn=2,
                 priors_table=None):
    """R/bayes.R:126-173 -> N x 4 array with columns CNn, CNt, Mt, LPP (R/bayes.R:171)."""
    f = _lib.vec(mufreq)
    n = f.size
    pc, pv = _priors(priors_table, CNt_min, CNt_max)
    out = np.empty((4, n))
    L = _lib.lib()
    if (np.ndim(weight_mufreq) == 0 and np.ndim(weight_ratio) == 0 and
            L.sqz_mufreq_scalar_table_bytes(float(weight_mufreq), float(weight_ratio)) > 0):
        # scalar weights (the reference's defaults): table-driven streaming kernel
        r = _lib.vec(depth_ratio, n, "depth.ratio")
        _lib.check(L.sqz_mufreq_point_scalar(
            _lib.ptr(f), _lib.ptr(r), n, float(weight_mufreq), float(weight_ratio),
            float(cellularity), float(ploidy), float(avg_depth_ratio), int(CNt_min),
            int(CNt_max), int(CNn), _lib.ptr(pc), _lib.ptr(pv), pc.size, _lib.ptr(out)))
        return out.T.copy()
    v = [_lib.vec(a, n, nm) for a, nm in ((f, "mufreq"), (depth_ratio, "depth.ratio"),
                                          (weight_mufreq, "weight.mufreq"),
                                          (weight_ratio, "weight.ratio"))]
    _lib.check(L.sqz_mufreq_point(
        *[_lib.ptr(a) for a in v], n, float(cellularity), float(ploidy),
        float(avg_depth_ratio), int(CNt_min), int(CNt_max), int(CNn), _lib.ptr(pc),
        _lib.ptr(pv), pc.size, _lib.ptr(out)))
    return out.T.copy()


def mufreq_grid_sums(mufreq, depth_ratio, cellularity, ploidy, avg_depth_ratio,
                     weight_mufreq=100, weight_ratio=100, CNt_min=1, CNt_max=7, CNn=2,
                     priors_table=None, n_gpus=1):
    """result$LPP of R/bayes.R:186 (ploidy fastest), from sqz_mufreq_grid."""
    f = _lib.vec(mufreq)
    n = f.size
    v = [_lib.vec(a, n, nm) for a, nm in ((f, "mufreq"), (depth_ratio, "depth.ratio"),
                                          (weight_mufreq, "weight.mufreq"),
                                          (weight_ratio, "weight.ratio"))]
    pl, ce = _lib.vec(ploidy), _lib.vec(cellularity)
    pc, pv = _priors(priors_table, CNt_min, CNt_max)
    out = np.empty(pl.size * ce.size)
    _lib.check(_lib.lib().sqz_mufreq_grid(
        *[_lib.ptr(a) for a in v], n, _lib.ptr(pl), pl.size, _lib.ptr(ce), ce.size,
        float(avg_depth_ratio), int(CNt_min), int(CNt_max), int(CNn), _lib.ptr(pc),
        _lib.ptr(pv), pc.size, int(n_gpus), _lib.ptr(out)))
    return out


def mufreq_model_fit(cellularity=None, ploidy=None, mc_cores=2, n_gpus=None, **kwargs):
    """R/bayes.R:175-194."""
    cellularity = r_seq(0.3, 1, 0.01) if cellularity is None else cellularity
    ploidy = r_seq(1, 7, 0.1) if ploidy is None else ploidy
    if n_gpus is None:
        n_gpus = max(1, _lib.lib().sqz_device_count())
    sums = mufreq_grid_sums(cellularity=cellularity, ploidy=ploidy, n_gpus=n_gpus, **kwargs)
    return finish_model_fit(sums, ploidy, cellularity)
