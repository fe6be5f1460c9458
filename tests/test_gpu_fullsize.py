"""Full-size parity of the BASELINE configurations against the CPU oracle, through the
kernels that actually run them (R/bayes.R:109-113: sum(baf.bayes(...)[, 4]) per grid point).

The oracle needs ~0.2 s of CPU per grid point at these segment counts, so it is run on
windows of the grid (>= 128 points in all, chosen to include the fitted solution) while
the GPU computes the whole grid.  The kernel path is asserted through the library's own
counters, so a planner change cannot silently move these tests to another kernel.

Also here: the bits of a grid point must not depend on how the grid is sharded (column
blocks or unit ranges for 1, 2, 4 and 8 GPUs), which makes the selected (cellularity,
ploidy) the same for any GPU count by construction; the top-2 margin of the fit is
reported and checked against the LPP tolerance."""
import ctypes as C
import os

import numpy as np
import pytest

import sequenza_b200 as S
from sequenza_b200 import _lib
from sequenza_b200.synth import DEFAULT_SEED, baf_args, synth_segments

pytestmark = pytest.mark.gpu
RTOL = 1e-9
PRI = {"CN": [2], "value": [2]}  # sequenza.fit's default priors.table (R/fit.R:6)
CORES = os.cpu_count() or 1


def grid_with_path(fn):
    """Run fn() with the library's per-call counters on; returns (result, kernel path):
    0 fused, 1 split, 2 pruned (segment per lane), 3 pruned (column per lane)."""
    import torch
    L = _lib.lib()
    L.sqz_profile_enable(1)
    try:
        out = fn()
        torch.cuda.synchronize()
        path, n_eval, launches = C.c_int(-1), C.c_ulonglong(0), C.c_longlong(0)
        L.sqz_profile_counters(C.byref(path), C.byref(n_eval), C.byref(launches))
        assert launches.value >= 4
    finally:
        L.sqz_profile_enable(0)
    return out, path.value


def oracle_windows(oracle, windows, **kw):
    """{(g0, g1): oracle sums} for linear windows of the grid (ploidy fastest)."""
    return {(g0, g1): oracle.baf_grid_sums(g0=g0, g1=g1, nthreads=CORES, **kw)
            for g0, g1 in windows}


def check_windows(got, ref_windows):
    worst = 0.0
    for (g0, g1), ref in ref_windows.items():
        g = got[g0:g1]
        assert np.isfinite(ref).all()
        rel = np.max(np.abs(g - ref) / np.abs(ref))
        worst = max(worst, rel)
        assert rel <= RTOL, ((g0, g1), rel)
        assert int(np.argmax(g)) == int(np.argmax(ref)), (g0, g1)
    return worst


def top2_margin(sums):
    """Relative gap between the best and the second-best grid point."""
    k = int(np.argmax(sums))
    rest = np.delete(sums, k)
    return k, (sums[k] - rest.max()) / abs(sums[k])


def windows_around(k, nP, G, width=32, extra=()):
    """A window of `width` grid points around linear index k (same column), plus others."""
    ic, ip = divmod(k, nP)
    lo = max(0, min(ip - width // 2, nP - width))
    w = [(ic * nP + lo, ic * nP + lo + width)]
    for g0 in extra:
        g0 = max(0, min(g0, G - width))
        w.append((g0, g0 + width))
    return w


def test_config2_full_size_vs_oracle(oracle):
    """BASELINE configs[1] exactly as bench.py runs it: 3e4 segments, CNt.max = 20, default
    91 x 61 grid -> column-per-lane kernel with 8-wide tiles.  128 oracle grid points at
    full N; the split kernel (forced) is held to the same windows."""
    seg = synth_segments(30000, seed=DEFAULT_SEED)
    ce, pl = S.r_seq(0.1, 1, 0.01), S.r_seq(1, 7, 0.1)
    kw = dict(cellularity=ce, ploidy=pl, CNt_max=20, priors_table=PRI, **baf_args(seg))
    got, path = grid_with_path(lambda: S.baf_grid_sums(**kw))
    assert path == 3, "config 2 is expected on the column-per-lane kernel, got path %d" % path
    k, margin = top2_margin(got)
    nP, G = pl.size, got.size
    wins = windows_around(k, nP, G, extra=(0, 45 * nP + 10, G - 32))  # incl. cellularity = 1
    refs = oracle_windows(oracle, wins, **kw)
    worst = check_windows(got, refs)
    os.environ["SQZ_COLWARP"] = "0"
    try:
        split, path = grid_with_path(lambda: S.baf_grid_sums(**kw))
    finally:
        del os.environ["SQZ_COLWARP"]
    assert path == 1
    check_windows(split, refs)
    assert int(np.argmax(split)) == k
    ref = oracle.baf_grid_sums(g0=wins[0][0], g1=wins[0][1], nthreads=CORES, **kw)
    assert wins[0][0] + int(np.argmax(ref)) == k  # the selected solution, bit-exact
    assert margin > 10 * RTOL, "top-2 margin %.3g too small for the tolerance to decide" % margin
    print("config 2: max rel err %.2e on 128 oracle points; top-2 margin %.3e" % (worst, margin))


def test_config3_full_size_vs_oracle(oracle):
    """BASELINE configs[2]: 1e5 segments, CNt.max = 7, the whole 901 x 601 grid -> the
    column-per-lane kernel with 32-wide tiles and 128 segment chunks.  160 oracle grid
    points at full N, the window around the selected solution among them."""
    import torch
    from sequenza_b200.dist import BafGridDevice, grid_plan, shard_columns, shard_units
    seg = synth_segments(100000, seed=DEFAULT_SEED)
    ce, pl = S.r_seq(0.1, 1, 0.001), S.r_seq(1, 7, 0.01)
    nC, nP = ce.size, pl.size
    plan = grid_plan(100000, 1, pl, nC)
    assert (plan["family"], plan["W"], plan["PT"]) == (3, 32, 32)
    dev = BafGridDevice(**baf_args(seg))
    full, path = grid_with_path(lambda: dev.run(pl, ce, priors_table=PRI))
    assert path == 3, "config 3 is expected on the column-per-lane kernel, got %d" % path
    got = full[0].cpu().numpy()
    assert np.isfinite(got).all()
    k, margin = top2_margin(got)
    G = got.size
    wins = windows_around(k, nP, G, extra=(0, 300 * nP + 17, 569 * nP + 580, G - 32))
    kw = dict(cellularity=ce, ploidy=pl, priors_table=PRI, **baf_args(seg))
    refs = oracle_windows(oracle, wins, **kw)
    worst = check_windows(got, refs)
    assert wins[0][0] + int(np.argmax(refs[wins[0]])) == k
    assert abs(ce[k // nP] - 0.7) <= 0.005 and abs(pl[k % nP] - 2.4) <= 0.02
    assert margin > 10 * RTOL, "top-2 margin %.3g too small for the tolerance to decide" % margin
    print("config 3: max rel err %.2e on 160 oracle points; solution c=%.3f p=%.2f; "
          "top-2 margin %.3e" % (worst, ce[k // nP], pl[k % nP], margin))

    # the same bits from any shard: unit ranges of 2, 4 and 8 GPUs ...
    for world in (2, 4, 8):
        buf = torch.zeros_like(full)
        for r in range(world):
            u0, u1 = shard_units(plan["units"], world, r)
            dev.run_units(pl, ce, u0, u1, out=buf, priors_table=PRI)
        assert torch.equal(buf, full), "unit shards of %d GPUs differ from the whole grid" % world
        assert int(torch.argmax(buf)) == k
    # ... and column blocks (first and last rank of an 8-way split: 112 and 113 columns)
    for r in (0, 7):
        c0, c1 = shard_columns(nC, 8, r)
        part = dev.run(pl, ce, c0, c1, priors_table=PRI)
        assert torch.equal(part[0], full[0, c0 * nP:c1 * nP]), (r, c0, c1)


def test_config3_ratio_priority_full_size_vs_oracle(oracle):
    """ratio.priority = TRUE (R/bayes.R:80-89) at 1e5 segments on a block of the fine grid."""
    seg = synth_segments(100000, seed=DEFAULT_SEED + 1)
    ce, pl = S.r_seq(0.64, 0.76, 0.001), S.r_seq(1, 7, 0.01)
    kw = dict(cellularity=ce, ploidy=pl, priors_table=PRI, ratio_priority=True,
              **baf_args(seg))
    got, path = grid_with_path(lambda: S.baf_grid_sums(**kw))
    assert path == 1  # ratio.priority runs the split kernel
    k, margin = top2_margin(got)
    wins = windows_around(k, pl.size, got.size, extra=(0, 60 * pl.size + 300, got.size - 32))
    worst = check_windows(got, oracle_windows(oracle, wins, **kw))
    assert margin > 10 * RTOL
    print("config 3, ratio.priority: max rel err %.2e; top-2 margin %.3e" % (worst, margin))


def test_ratio_priority_near_tie_goes_to_the_faithful_kernel(oracle):
    """A segment whose depth ratio sits half-way (in log) between the models of two copy
    numbers: (1 + x^2) of the two differ by rounding only, and the reference decides on
    dt() + log(prior).  The fast kernel must not decide such a case itself: the grid
    point is flagged and recomputed, so the sums equal the oracle's although the BAF
    terms of the two copy numbers differ by far more than rounding."""
    from sequenza_b200.model import theoretical_depth_ratio
    ce, pl = np.array([0.55, 0.8]), S.r_seq(1.5, 3.0, 0.1)
    seg = synth_segments(300, seed=5, cellularity=0.55, ploidy=2.0)
    c, p = ce[0], pl[4]
    for i, (a, b) in enumerate(((2, 3), (1, 2), (3, 4), (4, 5))):
        ra = theoretical_depth_ratio(a, c, p)
        rb = theoretical_depth_ratio(b, c, p)
        seg["depth_ratio"][10 * i] = np.sqrt(ra * rb)
        seg["Bf"][10 * i] = 0.33 + 0.03 * i
    kw = dict(cellularity=ce, ploidy=pl, ratio_priority=True, **baf_args(seg))  # flat priors
    got = S.baf_grid_sums(**kw)
    ref = oracle.baf_grid_sums(**kw)
    assert np.max(np.abs(got - ref) / np.abs(ref)) <= RTOL
    assert int(np.argmax(got)) == int(np.argmax(ref))
    slow = S.baf_grid_sums(mode=S.SQZ_MODE_FAITHFUL, **kw)
    np.testing.assert_array_equal(got[4], slow[4])  # the tied grid point took the repair
