"""segment.breaks (R/segments.R:113-198): the per-segment statistics that turn the seqz rows
and the breakpoints of a genome into the segment table sequenza.fit consumes.  The
breakpoints themselves (copynumber::aspcf, R/segments.R:1-111) stay with the caller.

The reductions and the kernel-density mode of b_allele_freq (R/model.R:69-98) run on the
GPU (sqz_segment_stats); the host keeps what the reference does around them: unique.breaks,
the every-other-interval selection, chromosome order and the N.ratio / length filter.
"""
from __future__ import annotations

import numpy as np

from . import _lib

SEGMENT_COLUMNS = ("Bf", "N.BAF", "sd.BAF", "depth.ratio", "N.ratio", "sd.ratio")


def unique_breaks(b, offset=1.0):
    """R/segments.R:133-138: bump repeated break positions until all differ, then sort as
    cut() does (cut.default: breaks <- sort.int(as.double(breaks)))."""
    b = np.array(b, dtype=np.float64)
    while True:
        d = np.diff(b) == 0
        if not d.any():
            break
        idx = np.flatnonzero(d)
        b[idx + 1] = b[idx] + offset
    return np.sort(b)


def layout_genome(chromosome, position, breaks):
    """Order the rows by chromosome (first appearance, R/segments.R:125 chr.order) and
    position, and build the offsets the kernels want.  `breaks` is a dict with keys chrom,
    start.pos, end.pos (R/segments.R:108).  Returns (row order, chromosome names, row_off,
    seg_off, breaks_vect)."""
    chromosome = np.asarray(chromosome)
    position = np.asarray(position, dtype=np.float64)
    names, first = np.unique(chromosome, return_index=True)
    names = names[np.argsort(first)]
    order, row_off, seg_off, brk = [], [0], [0], []
    bchrom = np.asarray(breaks["chrom"])
    for nm in names:
        rows = np.flatnonzero(chromosome == nm)
        rows = rows[np.argsort(position[rows], kind="stable")]
        order.append(rows)
        row_off.append(row_off[-1] + rows.size)
        sel = np.flatnonzero(bchrom == nm)
        inter = np.empty(2 * sel.size)
        inter[0::2] = np.asarray(breaks["start.pos"], dtype=np.float64)[sel]
        inter[1::2] = np.asarray(breaks["end.pos"], dtype=np.float64)[sel]
        brk.append(unique_breaks(inter))
        seg_off.append(seg_off[-1] + sel.size)
    return (np.concatenate(order) if order else np.empty(0, dtype=np.int64), names,
            np.asarray(row_off, dtype=np.int64), np.asarray(seg_off, dtype=np.int64),
            np.concatenate(brk) if brk else np.empty(0))


def segment_stats(row_off, seg_off, position, adjusted_ratio, depth_normal, good_reads, Af, Bf,
                  is_het, breaks_vect, min_reads_baf=1, weighted_mean=True):
    """sqz_segment_stats: n_seg x 6 array (columns SEGMENT_COLUMNS) for a genome laid out as
    concatenated, position-sorted chromosomes."""
    ro = np.ascontiguousarray(row_off, dtype=np.int64)
    so = np.ascontiguousarray(seg_off, dtype=np.int64)
    n = int(ro[-1]) if ro.size else 0
    cols = [_lib.vec(v, n) for v in (position, adjusted_ratio, depth_normal, good_reads, Af, Bf)]
    het = np.ascontiguousarray(is_het, dtype=np.uint8)
    brk = _lib.vec(breaks_vect)
    S = int(so[-1]) if so.size else 0
    out = np.empty((6, S))
    _lib.check(_lib.lib().sqz_segment_stats(
        max(ro.size - 1, 0), _lib.ptr(ro), _lib.ptr(so), *[_lib.ptr(c) for c in cols],
        _lib.ptr(het), _lib.ptr(brk), float(min_reads_baf), int(bool(weighted_mean)),
        _lib.ptr(out)))
    return out.T.copy()


def segment_breaks(seqz_tab, breaks, min_reads_baf=1, weighted_mean=True):
    """R/segments.R:113-198.  seqz_tab: dict of columns chromosome, position,
    zygosity.normal, good.reads, Af, Bf, adjusted.ratio, depth.normal; breaks: dict chrom,
    start.pos, end.pos.  Returns a dict of columns chromosome, start.pos, end.pos, Bf,
    N.BAF, sd.BAF, depth.ratio, N.ratio, sd.ratio, filtered as the reference does
    (N.ratio per Mb >= 2, R/segments.R:196-197)."""
    order, names, row_off, seg_off, brk = layout_genome(seqz_tab["chromosome"],
                                                        seqz_tab["position"], breaks)
    col = lambda k: np.asarray(seqz_tab[k], dtype=np.float64)[order]  # noqa: E731
    het = (np.asarray(seqz_tab["zygosity.normal"])[order] == "het")
    st = segment_stats(row_off, seg_off, col("position"), col("adjusted.ratio"),
                       col("depth.normal"), col("good.reads"), col("Af"), col("Bf"), het, brk,
                       min_reads_baf, weighted_mean)
    chrom = np.repeat(names, np.diff(seg_off))
    start, end = brk[0::2], brk[1::2]
    keep = (st[:, 4] / ((end - start) / 1e6)) >= 2
    out = {"chromosome": chrom[keep], "start.pos": start[keep], "end.pos": end[keep]}
    for k, name in enumerate(SEGMENT_COLUMNS):
        out[name] = st[keep, k]
    return out


def synth_seqz(n_chr=3, rows_per_chr=4000, segs_per_chr=7, seed=1, het_frac=0.3, depth=60):
    """A synthetic genome in the layout segment_stats takes: positions sorted within each
    chromosome, piecewise-constant depth ratio and BAF per segment with noise, some NA
    ratios, and breakpoints that sit exactly on positions (to exercise cut()'s half-open
    intervals) with a gap between consecutive segments."""
    rng = np.random.Generator(np.random.PCG64(seed))
    pos, ratio, dn, gr, af, bf, het, brk = [], [], [], [], [], [], [], []
    row_off, seg_off = [0], [0]
    for _ in range(n_chr):
        cb = []
        p = np.cumsum(rng.integers(1, 80, rows_per_chr))  # sorted, distinct positions
        cuts = np.sort(rng.choice(np.arange(5, rows_per_chr - 5), segs_per_chr - 1,
                                  replace=False))
        edges = np.concatenate([[0], cuts, [rows_per_chr - 1]])
        r = np.empty(rows_per_chr)
        b = np.empty(rows_per_chr)
        g = np.round(rng.uniform(0.5 * depth, 1.5 * depth, rows_per_chr))
        for k in range(segs_per_chr):
            lo, hi = edges[k], edges[k + 1] + 1
            r[lo:hi] = rng.choice([0.6, 0.8, 1.0, 1.25, 1.5]) * np.exp(rng.normal(0, 0.2, hi - lo))
            f = rng.choice([0.1, 0.25, 0.33, 0.5])
            bb = rng.binomial(g[lo:hi].astype(int), f) / g[lo:hi]
            b[lo:hi] = np.minimum(bb, 1 - bb)
            # (start, end]: the start is a real position (excluded), the end too (included)
            cb += [float(p[lo]), float(p[edges[k + 1]] - (1 if k < segs_per_chr - 1 else 0))]
        brk += list(unique_breaks(cb))
        r[rng.random(rows_per_chr) < 0.01] = np.nan
        pos.append(p.astype(np.float64))
        ratio.append(r)
        dn.append(np.round(rng.uniform(10, 90, rows_per_chr)))
        gr.append(g)
        bf.append(b)
        af.append(1 - b)
        het.append((rng.random(rows_per_chr) < het_frac).astype(np.uint8))
        row_off.append(row_off[-1] + rows_per_chr)
        seg_off.append(seg_off[-1] + segs_per_chr)
    cat = np.concatenate
    return {"row_off": np.asarray(row_off, dtype=np.int64),
            "seg_off": np.asarray(seg_off, dtype=np.int64), "position": cat(pos),
            "adjusted_ratio": cat(ratio), "depth_normal": cat(dn), "good_reads": cat(gr),
            "Af": cat(af), "Bf": cat(bf), "is_het": cat(het),
            "breaks_vect": np.asarray(brk, dtype=np.float64)}
