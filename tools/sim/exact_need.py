#!/usr/bin/env python3
"""Simulation (numpy, no GPU): candidates per (segment, tile) of the column-per-lane kernel
under its rule (bound over the tile < Qmax of the first candidate) against the smallest
set any rule could reach (copy numbers that are the minimum for at least one ploidy of the
tile), warp-level (longest list among the 32 lanes)."""
import sys
import numpy as np
sys.path.insert(0, ".")
from sequenza_b200.synth import synth_segments
from sequenza_b200.model import theoretical_depth_ratio, theoretical_baf

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
PT = 32
seg = synth_segments(100000)
idx = np.random.default_rng(0).choice(100000, N, replace=False)
seg = {k: (np.asarray(v)[idx] if np.ndim(v) == 1 and len(v) == 100000 else v) for k, v in seg.items()}
ce_all = np.round(np.arange(0.1, 1.0005, 0.001), 3)
pl_all = np.round(np.arange(1, 7.005, 0.01), 2)
cn = np.arange(0, 8)
pr = np.ones(8); pr[2] = 2.0
rows = cn // 2 + 1
pr = pr / (pr * rows).sum()
w = pr ** (-1 / 3)
kr = 1 / (seg["sd_ratio"] / np.sqrt(seg["weight_ratio"])) / np.sqrt(5)
lr = np.log(seg["depth_ratio"])
Bf = seg["Bf"]; kb = 1 / (seg["sd_Bf"] / np.sqrt(seg["weight_Bf"])) / np.sqrt(5)
hasB = ~np.isnan(Bf)
nt = (len(pl_all) + PT - 1) // PT
tiles = list(range(0, nt, 2))
tot = {"rule": [], "exact": [], "mid": [], "h2": [], "h4": []}
for c0 in (0, 160, 320, 480, 640, 800):
    R = np.zeros((32, len(tiles), N), dtype=np.int8)
    E = np.zeros_like(R); M = np.zeros_like(R); H2 = np.zeros_like(R); H4 = np.zeros_like(R)
    for a in range(32):
        cell = ce_all[c0 + a]
        wub = np.empty((8, N))
        for c in cn:
            mb = np.array([theoretical_baf(c, b, cell) for b in range(c // 2 + 1)])
            xb = (mb[:, None] - Bf[None, :]) * kb[None, :]
            m0 = np.where(hasB, np.nanmin(np.abs(np.where(hasB, xb, 0.0)), axis=0), 0.0)
            wub[c] = w[c] * (1 + m0 * m0)
        for b, it in enumerate(tiles):
            pj = pl_all[it * PT:(it + 1) * PT]
            if len(pj) < PT:
                pj = np.concatenate([pj, np.repeat(pj[-1], PT - len(pj))])
            l = np.log(theoretical_depth_ratio(cn[:, None], cell, pj[None, :]))
            x = (l[:, :, None] - lr[None, None, :]) * kr[None, None, :]
            q = wub[:, None, :] * (1 + x * x)                                   # [8, pt, N]
            xf, xl = x[:, 0, :], x[:, -1, :]
            tmin = np.where(xf * xl <= 0, 0.0, np.minimum(np.abs(xf), np.abs(xl)))
            LB = wub * (1 + tmin * tmin)
            first = np.argmin(LB, axis=0)
            qf = np.take_along_axis(q, first[None, None, :].repeat(PT, 1), 0)[0]  # [pt, N]
            need = LB < qf.max(axis=0)[None, :]
            need[first, np.arange(N)] = True
            R[a, b] = need.sum(axis=0)
            am = np.argmin(q, axis=0)                                           # [pt, N]
            ex = np.zeros((8, N), dtype=bool)
            for c in cn:
                ex[c] = (am == c).any(axis=0)
            E[a, b] = ex.sum(axis=0)
            # middle: exact comparison against the first candidate only
            md = (q < qf[None, :, :]).any(axis=1)
            md[first, np.arange(N)] = True
            M[a, b] = md.sum(axis=0)
            for parts, H in ((2, H2), (4, H4)):
                nd = np.zeros((8, N), dtype=bool)
                w_ = PT // parts
                for h in range(parts):
                    xa, xz = x[:, h * w_, :], x[:, min((h + 1) * w_, PT - 1), :]
                    tm = np.where(xa * xz <= 0, 0.0, np.minimum(np.abs(xa), np.abs(xz)))
                    lbh = wub * (1 + tm * tm)
                    qe = np.maximum(qf[h * w_], qf[min((h + 1) * w_, PT - 1)])
                    nd |= lbh < qe[None, :]
                nd[first, np.arange(N)] = True
                H[a, b] = nd.sum(axis=0)
    r, e, m_ = R.max(axis=0).mean(), E.max(axis=0).mean(), M.max(axis=0).mean()
    h = np.bincount(R.max(axis=0).ravel(), minlength=9)[:9] / R.max(axis=0).size
    he = np.bincount(E.max(axis=0).ravel(), minlength=9)[:9] / R.max(axis=0).size
    print("cols %3d..: rule %.3f | beats-first %.3f | exact %.3f   hist rule %s exact %s" % (
        c0, r, m_, e, np.round(h[1:5], 3), np.round(he[1:5], 3)), flush=True)
    tot["rule"].append(r); tot["exact"].append(e); tot["mid"].append(m_)
    tot["h2"].append(H2.max(axis=0).mean()); tot["h4"].append(H4.max(axis=0).mean())
    print("           halves %.3f quarters %.3f  hist halves %s" % (tot["h2"][-1], tot["h4"][-1], np.round(np.bincount(H2.max(axis=0).ravel(), minlength=9)[1:5] / H2.max(axis=0).size, 3)))
print("mean: rule %.3f, beats-first %.3f, exact %.3f, halves %.3f, quarters %.3f" % tuple(np.mean(tot[k]) for k in ("rule", "mid", "exact", "h2", "h4")))
