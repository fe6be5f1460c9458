#!/usr/bin/env python3
"""Simulation (numpy, no GPU): how many copy-number evaluations a WARP of the pruned grid
kernel performs per segment under different lane mappings.  Candidate rule as in
baf_grid_pruned_kernel (exact interval bound form)."""
import sys
import numpy as np
sys.path.insert(0, ".")
from sequenza_b200.synth import synth_segments
from sequenza_b200.model import theoretical_depth_ratio, theoretical_baf

N = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
PT = int(sys.argv[2]) if len(sys.argv) > 2 else 32
seg = synth_segments(N)
ce_all = np.round(np.arange(0.1, 1.0005, 0.001), 3)
pl_all = np.round(np.arange(1, 7.005, 0.01), 2)
cn = np.arange(0, 8)
w = np.ones(8); w[2] = 2.0 ** (-1 / 3)
kr = 1 / (seg["sd_ratio"] / np.sqrt(seg["weight_ratio"])) / np.sqrt(5)
lr = np.log(seg["depth_ratio"])
Bf = seg["Bf"]; kb = 1 / (seg["sd_Bf"] / np.sqrt(seg["weight_Bf"])) / np.sqrt(5)
hasB = ~np.isnan(Bf)

def counts(cols, tiles):
    """need-count (incl. the first) [col, tile, seg]"""
    out = np.zeros((len(cols), len(tiles), N), dtype=np.int8)
    for a, ic in enumerate(cols):
        cell = ce_all[ic]
        wub = np.empty((8, N))
        for c in cn:
            nb = c // 2 + 1
            mb = np.array([theoretical_baf(c, b, cell) for b in range(nb)])
            xb = (mb[:, None] - Bf[None, :]) * kb[None, :]
            m0 = np.where(hasB, np.nanmin(np.abs(np.where(hasB, xb, 0.0)), axis=0), 0.0)
            wub[c] = w[c] * (1 + m0 * m0)
        for b, it in enumerate(tiles):
            pj = pl_all[it * PT:(it + 1) * PT]
            l = np.log(theoretical_depth_ratio(cn[:, None], cell, pj[None, :]))  # [8, pt]
            x = (l[:, :, None] - lr[None, None, :]) * kr[None, None, :]        # [8, pt, N]
            q = wub[:, None, :] * (1 + x * x)
            xf, xl = x[:, 0, :], x[:, -1, :]
            tmin = np.where(xf * xl <= 0, 0.0, np.minimum(np.abs(xf), np.abs(xl)))
            LB = wub * (1 + tmin * tmin)
            first = np.argmin(LB, axis=0)
            qf = np.take_along_axis(q, first[None, None, :].repeat(q.shape[1], 1), 0)[0]
            qmax = qf.max(axis=0)
            need = LB < qmax[None, :]
            need[first, np.arange(N)] = True
            out[a, b] = need.sum(axis=0)
    return out

nt = (len(pl_all) + PT - 1) // PT
tiles = list(range(0, nt, 3))
res = []
for c0 in (0, 200, 450, 700, 869):
    cnt = counts(range(c0, c0 + 32), tiles).astype(float)  # [32, T, N]
    per_thread = cnt.mean()
    A = cnt.reshape(32, len(tiles), N // 32, 32).max(axis=3).mean()
    B = cnt.max(axis=0).mean()
    B16 = cnt.reshape(2, 16, len(tiles), N).max(axis=1).mean()
    res.append((c0, per_thread, A, B, B16))
    print("cols %3d..: per-thread %.2f | warp = 32 segments %.2f | warp = 32 columns %.2f | 16 columns x 2 segs %.2f"
          % (c0, per_thread, A, B, B16), flush=True)
r = np.array(res)
print("mean: per-thread %.2f, A %.2f, B %.2f" % (r[:, 1].mean(), r[:, 2].mean(), r[:, 3].mean()))
