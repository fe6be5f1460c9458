"""sequenza.fit (R/fit.R:1-69): segment / mutation filtering and dispatch, host code as
in the reference; the grid search itself runs on the GPU (bayes.baf_model_fit)."""
from __future__ import annotations

import numpy as np

from .bayes import baf_model_fit, mufreq_model_fit, r_seq

_SEG_COLS = ("chromosome", "start.pos", "end.pos", "Bf", "N.BAF", "sd.BAF", "depth.ratio",
             "N.ratio", "sd.ratio")


def _columns(tab, cols):
    """One per-chromosome table (dict of arrays or pandas.DataFrame) -> dict of arrays."""
    if tab is None:
        return None
    out = {}
    for c in cols:
        v = tab[c]
        v = v.to_numpy() if hasattr(v, "to_numpy") else np.asarray(v)
        out[c] = v.astype(str) if c == "chromosome" else v.astype(np.float64)
    return out


def _rbind(tables, cols, chromosome_list):
    """do.call(rbind, x[chromosome.list]): positional, 1-based, missing entries skipped."""
    if isinstance(tables, dict):
        tables = list(tables.values())
    if chromosome_list is None:
        picked = list(tables)
    else:
        picked = [tables[i - 1] for i in chromosome_list if 1 <= i <= len(tables)]
    picked = [_columns(t, cols) for t in picked if t is not None]
    if not picked:
        return {c: np.empty(0, dtype=str if c == "chromosome" else np.float64) for c in cols}
    return {c: np.concatenate([t[c] for t in picked]) for c in cols}


def sequenza_fit(sequenza_extract, female=True, N_ratio_filter=10, N_BAF_filter=1,
                 segment_filter=3e6, mufreq_treshold=0.10, XY=None, cellularity=None,
                 ploidy=None, ratio_priority=False, method="baf", priors_table=None,
                 chromosome_list=tuple(range(1, 25)), mc_cores=2, n_gpus=None):
    """R/fit.R:1-69.  `sequenza_extract` is a mapping with `segments` (and, for
    method="mufreq", `mutations`): ordered per-chromosome tables, and `avg.depth.ratio`.
    Returns list(ploidy, cellularity, lpp) as a dict, ready for get_ci()."""
    XY = {"X": "X", "Y": "Y"} if XY is None else XY
    cellularity = r_seq(0.1, 1, 0.01) if cellularity is None else cellularity
    ploidy = r_seq(1, 7, 0.1) if ploidy is None else ploidy
    priors_table = {"CN": [2], "value": [2]} if priors_table is None else priors_table
    avg_depth_ratio = sequenza_extract["avg.depth.ratio"]

    if method == "baf":
        segs = _rbind(sequenza_extract["segments"], _SEG_COLS, chromosome_list)
        segs_len = segs["end.pos"] - segs["start.pos"]
        sd_baf, sd_ratio = segs["sd.BAF"].copy(), segs["sd.ratio"].copy()
        if sd_baf.size:  # R/fit.R:29-32: zero sd -> the largest sd
            sd_baf[sd_baf == 0] = np.nanmax(sd_baf)
            sd_ratio[sd_ratio == 0] = np.nanmax(sd_ratio)
        with np.errstate(invalid="ignore"):
            filt = (segs["N.ratio"] > N_ratio_filter) & (segs["N.BAF"] > N_BAF_filter)
            filt = (segs_len >= segment_filter) & filt
        if female:
            is_xy = segs["chromosome"] == XY["Y"]
        else:
            is_xy = np.isin(segs["chromosome"], list(XY.values()))
        keep = filt & ~is_xy
        seg_len_mb = segs_len[keep] / 1e6
        return baf_model_fit(
            Bf=segs["Bf"][keep], depth_ratio=segs["depth.ratio"][keep],
            sd_ratio=sd_ratio[keep], weight_ratio=seg_len_mb, sd_Bf=sd_baf[keep],
            weight_Bf=seg_len_mb, avg_depth_ratio=avg_depth_ratio, cellularity=cellularity,
            ploidy=ploidy, priors_table=priors_table, mc_cores=mc_cores,
            ratio_priority=ratio_priority, n_gpus=n_gpus)
    elif method == "mufreq":
        cols = ("chromosome", "F", "good.reads", "adjusted.ratio")
        mut = _rbind(sequenza_extract["mutations"], cols, chromosome_list)
        ok = ~(np.isnan(mut["F"]) | np.isnan(mut["good.reads"]) |
               np.isnan(mut["adjusted.ratio"]))  # na.exclude
        mut = {c: v[ok] for c, v in mut.items()}
        filt = mut["F"] >= mufreq_treshold
        if female:
            is_xy = mut["chromosome"] == XY["Y"]
        else:
            is_xy = np.isin(mut["chromosome"], list(XY.values()))
        keep = filt & ~is_xy
        w = np.rint(mut["good.reads"][keep])  # round(): half to even
        return mufreq_model_fit(
            mufreq=mut["F"][keep], depth_ratio=mut["adjusted.ratio"][keep],
            weight_ratio=2 * w, weight_mufreq=w, avg_depth_ratio=avg_depth_ratio,
            cellularity=cellularity, ploidy=ploidy, priors_table=priors_table,
            mc_cores=mc_cores, n_gpus=n_gpus)
    raise ValueError('The only available methods are "baf" and "mufreq"')
