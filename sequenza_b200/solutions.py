"""R/solutions.R: credible region, point estimate and alternative solutions from the
list(ploidy, cellularity, lpp) that the grid fit returns.  Host code, as in the
reference (O(G log G) on G doubles; SURVEY 8f ranks a device version as the next row)."""
from __future__ import annotations

import numpy as np


def _which_max(v):
    """R's which.max: first maximum, NaN skipped; -1 if nothing is comparable."""
    v = np.asarray(v, dtype=np.float64)
    ok = ~np.isnan(v)
    if not ok.any():
        return -1
    return int(np.flatnonzero(ok & (v == v[ok].max()))[0])


def get_ci(cp_table, level=0.95):
    """R/solutions.R:1-26.  lpp rows = ploidy, columns = cellularity."""
    lpp = np.asarray(cp_table["lpp"], dtype=np.float64)
    ploidy = np.asarray(cp_table["ploidy"], dtype=np.float64)
    cell = np.asarray(cp_table["cellularity"], dtype=np.float64)
    cache = getattr(cp_table, "ci_cache", None)
    if cache is not None and cache["lpp_id"] == id(cp_table["lpp"]) and np.ndim(level) == 0:
        # reductions already made on the device (sqz_fit_finish): a marginal maximum is at
        # or above the threshold element iff the mass strictly above it is < level
        vx, vy = cache["row_max"], cache["col_max"]
        sel_x = ploidy[cache["row_above"] < level]
        sel_y = cell[cache["col_above"] < level]
        if sel_x.size and sel_y.size:
            return {
                "values.ploidy": {"x": ploidy.copy(), "y": vx / np.sum(vx)},
                "confint.ploidy": (float(sel_x.min()), float(sel_x.max())),
                "max.ploidy": float(ploidy[_which_max(vx)]),
                "values.cellularity": {"x": vy / np.sum(vy), "y": cell.copy()},
                "confint.cellularity": (float(sel_y.min()), float(sel_y.max())),
                "max.cellularity": float(cell[_which_max(vy)]),
            }
    flat = lpp.reshape(-1)
    znormsort = np.sort(flat[~np.isnan(flat)])[::-1]          # sort(decreasing = TRUE)
    cum = np.cumsum(znormsort.astype(np.longdouble)).astype(np.float64)  # long double acc
    n = int(np.sum(cum < level))                              # sum(cum < x) + 1, 1-based
    thresh = znormsort[min(n, znormsort.size - 1)]
    with np.errstate(invalid="ignore"):
        vx = np.max(lpp, axis=1)                              # apply(lpp, 1, max)
        vy = np.max(lpp, axis=0)                              # apply(lpp, 2, max)
    sel_x = ploidy[vx >= thresh]
    sel_y = cell[vy >= thresh]
    return {
        "values.ploidy": {"x": ploidy.copy(), "y": vx / np.sum(vx)},
        "confint.ploidy": (float(sel_x.min()), float(sel_x.max())),
        "max.ploidy": float(ploidy[_which_max(vx)]),
        "values.cellularity": {"x": vy / np.sum(vy), "y": cell.copy()},
        "confint.cellularity": (float(sel_y.min()), float(sel_y.max())),
        "max.cellularity": float(cell[_which_max(vy)]),
    }


def alternative_cp_solutions(cp_table):
    """R/solutions.R:28-47 -> dict of columns cellularity, ploidy, SLPP (best first)."""
    ci = get_ci(cp_table)
    lpp = np.asarray(cp_table["lpp"], dtype=np.float64)
    ploidy = np.asarray(cp_table["ploidy"], dtype=np.float64)
    cell = np.asarray(cp_table["cellularity"], dtype=np.float64)
    y = ci["values.ploidy"]["y"]
    p_alt = np.flatnonzero(np.diff(np.sign(np.diff(y))) == -2) + 1
    rows = []
    for ip in p_alt:
        ic = _which_max(lpp[ip, :])
        rows.append((cell[ic], ploidy[ip], lpp[ip, ic]))
    if rows:
        rows.sort(key=lambda r: -r[2])  # order(SLPP, decreasing = TRUE); stable
    else:
        ip = np.flatnonzero(ploidy == ci["max.ploidy"])
        ic = np.flatnonzero(cell == ci["max.cellularity"])
        rows = [(ci["max.cellularity"], ci["max.ploidy"], float(lpp[ip[0], ic[0]]))]
    a = np.array(rows, dtype=np.float64).reshape(-1, 3)
    return {"cellularity": a[:, 0], "ploidy": a[:, 1], "SLPP": a[:, 2]}
