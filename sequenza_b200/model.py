"""Host-side helpers with the reference's names (R/model.R:1-67).

These are the small closed-form functions sequenza exports for plotting and for users;
they are NOT on the GPU path (the kernels build their own model tables on the device,
see csrc/sqz_common.cuh, and `device_model_points` returns those for checking).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib


def theoretical_depth_ratio(CNt, cellularity, ploidy, CNn=2, normal_ploidy=2,
                            avg_depth_ratio=1):
    """R/model.R:1-7."""
    CNt, cellularity, ploidy, CNn = map(np.asarray, (CNt, cellularity, ploidy, CNn))
    cellu_copy_term = (1 - cellularity) + (CNt / CNn * cellularity)
    ploidy_cellu_term = (ploidy / normal_ploidy * cellularity) + 1 - cellularity
    return avg_depth_ratio * cellu_copy_term / ploidy_cellu_term


def theoretical_baf(CNt, B, cellularity, CNn=2):
    """R/model.R:9-14; NA where CNn <= 1."""
    CNt, B, cellularity, CNn = np.broadcast_arrays(*map(
        lambda v: np.asarray(v, dtype=np.float64), (CNt, B, cellularity, CNn)))
    with np.errstate(divide="ignore", invalid="ignore"):
        baf = ((B * cellularity) + (1 - cellularity)) / \
              ((CNt * cellularity) + CNn * (1 - cellularity))
    baf = np.where(CNn <= 1, np.nan, baf)
    return baf if baf.ndim else float(baf)


def theoretical_mufreq(CNt, Mt, cellularity, CNn=2):
    """R/model.R:16-20."""
    CNt, Mt, cellularity, CNn = map(lambda v: np.asarray(v, dtype=np.float64),
                                    (CNt, Mt, cellularity, CNn))
    normal_alleles = (CNt - Mt) * cellularity + CNn * (1 - cellularity)
    all_alleles = (CNt * cellularity) + CNn * (1 - cellularity)
    return 1 - (normal_alleles / all_alleles)


def baf_types_matrix(CNt_min, CNt_max, CNn=2):
    """R/model.R:22-37 -> dict of columns CNn, CNt, B (rows: CNt asc, B asc)."""
    cnt, b = [], []
    for c in range(int(CNt_min), int(CNt_max) + 1):
        nb = 1 if CNn < 2 else c // 2 + 1
        cnt += [c] * nb
        b += list(range(nb))
    return {"CNn": np.full(len(cnt), float(CNn)), "CNt": np.array(cnt, dtype=np.float64),
            "B": np.array(b, dtype=np.float64)}


def mufreq_types_matrix(CNt_min, CNt_max, CNn=2):
    """R/model.R:39-47 -> dict of columns CNn, CNt, Mt (rows: CNt asc, Mt asc)."""
    cnt, mt = [], []
    for c in range(int(CNt_min), int(CNt_max) + 1):
        cnt += [c] * (c + 1)
        mt += list(range(c + 1))
    return {"CNn": np.full(len(cnt), float(CNn)), "CNt": np.array(cnt, dtype=np.float64),
            "Mt": np.array(mt, dtype=np.float64)}


def baf_model_points(cellularity, ploidy, baf_types, avg_depth_ratio):
    """R/model.R:49-56 -> dict BAF, depth.ratio."""
    return {
        "BAF": theoretical_baf(baf_types["CNt"], baf_types["B"], cellularity,
                               baf_types["CNn"]),
        "depth.ratio": theoretical_depth_ratio(baf_types["CNt"], cellularity, ploidy,
                                               baf_types["CNn"],
                                               avg_depth_ratio=avg_depth_ratio),
    }


def mufreq_model_points(cellularity, ploidy, mufreq_types, avg_depth_ratio):
    """R/model.R:58-67 -> dict mufreqs, depth.ratio."""
    return {
        "mufreqs": theoretical_mufreq(mufreq_types["CNt"], mufreq_types["Mt"], cellularity,
                                      mufreq_types["CNn"]),
        "depth.ratio": theoretical_depth_ratio(mufreq_types["CNt"], cellularity, ploidy,
                                               mufreq_types["CNn"],
                                               avg_depth_ratio=avg_depth_ratio),
    }


def device_model_points(kind, cellularity, ploidy, avg_depth_ratio=1.0, CNt_min=0,
                        CNt_max=7, CNn=2, priors_table=None):
    """Model table exactly as the CUDA kernels compute it (sqz_model_points).
    kind: "baf" or "mufreq".  Returns dict CNt, second, depth.ratio, value, prior."""
    from .bayes import _priors
    k = {"baf": 0, "mufreq": 1}[kind]
    pc, pv = _priors(priors_table, CNt_min, CNt_max)
    T = C.c_int(0)
    L = _lib.lib()
    _lib.check(L.sqz_model_points(k, cellularity, ploidy, avg_depth_ratio, CNt_min, CNt_max,
                                  CNn, _lib.ptr(pc), _lib.ptr(pv), pc.size, C.byref(T), None))
    out = np.empty((5, T.value))
    _lib.check(L.sqz_model_points(k, cellularity, ploidy, avg_depth_ratio, CNt_min, CNt_max,
                                  CNn, _lib.ptr(pc), _lib.ptr(pv), pc.size, C.byref(T),
                                  _lib.ptr(out)))
    return {"CNt": out[0], "second": out[1], "depth.ratio": out[2], "value": out[3],
            "prior": out[4]}
