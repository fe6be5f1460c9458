"""Launch planning and workspace sizing (host code inside the library; no GPU needed):
which kernel path and geometry `sqz_baf_grid_plan` reports for the BASELINE shapes, how the
grid is cut into units, and what `sqz_baf_grid_workspace*_bytes` ask for.  CPU only."""
import numpy as np
import pytest

from sequenza_b200 import _lib
from sequenza_b200.dist import grid_plan, shard_units, unit_points

FAMILY = {0: "fused", 1: "split", 2: "pruned", 3: "colwarp"}


def plan(n, s, nP, nC, cn_max, rp=0, ploidy=None):
    pl = np.linspace(1.0, 7.0, nP) if ploidy is None else ploidy
    return grid_plan(n, s, pl, nC, 0, cn_max, 2, bool(rp))


@pytest.fixture(autouse=True)
def no_overrides(monkeypatch):
    for k in ("SQZ_PT", "SQZ_THREADS", "SQZ_CHUNKS", "SQZ_SPLIT", "SQZ_PRUNE", "SQZ_COLWARP",
              "SQZ_CW_MAXCHUNKS"):
        monkeypatch.delenv(k, raising=False)


def test_config2_takes_narrow_column_per_lane_tiles():
    # 91 x 61 grid, CNt 0..20, 3e4 segments: 8-wide tiles, 3 column groups x 8 tiles = 24
    # units, and 74 chunks so that 24 x 74 one-warp CTAs are one wave of 148 SMs x 12
    p = plan(30000, 1, 61, 91, 20)
    assert (p["PT"], p["nPT"], p["threads"], p["T"]) == (8, 8, 32, 121)
    assert FAMILY[p["family"]] == "colwarp" and p["W"] == 32 and p["units"] == 24
    assert p["chunks"] == 74
    # a short table, ratio.priority or an unsorted ploidy axis: the split kernel
    q = plan(3000, 1, 61, 91, 20)
    assert (q["PT"], q["nPT"], q["threads"]) == (32, 2, 128) and FAMILY[q["family"]] == "split"
    assert q["W"] == 1 and q["units"] == 91 * 2 and 1 <= q["chunks"] <= 64
    assert FAMILY[plan(30000, 1, 61, 91, 20, rp=1)["family"]] == "split"


def test_config3_takes_the_column_per_lane_kernel():
    # one warp per CTA, 32-wide tiles, chunk count from the segment table alone
    p = plan(100000, 1, 601, 901, 7)
    assert (p["PT"], p["nPT"], p["threads"], p["T"]) == (32, 19, 32, 20)
    assert FAMILY[p["family"]] == "colwarp" and p["W"] == 32 and p["units"] == 29 * 19
    assert p["chunks"] == 128
    # the same grid over 3,000 segments (46 per chunk) takes 16-wide tiles
    q = plan(3000, 1, 601, 181, 7)
    assert (q["PT"], q["threads"]) == (16, 32)
    # a grid too narrow to fill the lanes of a 32-column group stays segment-per-lane
    assert plan(100000, 1, 601, 12, 7)["threads"] == 128
    # ratio.priority and an unsorted ploidy axis have no column-per-lane instantiation
    assert plan(100000, 1, 601, 901, 7, rp=1)["threads"] == 128
    shuffled = np.random.default_rng(0).permutation(np.linspace(1, 7, 601))
    assert FAMILY[plan(100000, 1, 601, 901, 7, ploidy=shuffled)["family"]] == "split"


def test_cohort_and_small_fits():
    # cohort: the samples fill the machine, few chunks (each costs a slab of partials)
    p = plan(512 * 3000, 512, 61, 91, 7)
    assert p["threads"] == 32 and p["PT"] == 16 and p["chunks"] == 4
    # the bundled example's size: segment-per-lane kernel
    assert plan(150, 1, 61, 91, 7)["threads"] in (64, 128)
    # CNt 0..20 on a fine ploidy axis: column-per-lane, 16-wide tiles
    p = plan(30000, 1, 601, 181, 20)
    assert (p["PT"], p["threads"]) == (16, 32)


def test_units_cover_the_grid_once_for_any_world_size():
    for nC, nP, n, cn in ((901, 601, 100000, 7), (91, 61, 30000, 20), (33, 40, 500, 7)):
        p = plan(n, 1, nP, nC, cn)
        for world in (1, 2, 3, 4, 8):
            seen = np.zeros((nC, nP), dtype=int)
            sizes = []
            for r in range(world):
                u0, u1 = shard_units(p["units"], world, r)
                sizes.append(u1 - u0)
                seen += unit_points(p, nC, nP, u0, u1)
            assert (seen == 1).all()
            assert max(sizes) - min(sizes) <= 1
    # config 3 on eight GPUs: 68 or 69 of the 551 units each
    p = plan(100000, 1, 601, 901, 7)
    assert sorted({shard_units(p["units"], 8, r)[1] - shard_units(p["units"], 8, r)[0]
                   for r in range(8)}) == [68, 69]


def test_workspace_sizes():
    L = _lib.lib()
    # minimum: grows with the grid, the segment count and the sample count
    base = L.sqz_baf_grid_workspace_bytes(30000, 1, 61, 91)
    assert base % 256 == 0 and base <= 16 << 20  # fits the no-alloc entry point's scratch
    assert L.sqz_baf_grid_workspace_bytes(30000, 1, 601, 901) > base
    assert L.sqz_baf_grid_workspace_bytes(300000, 1, 61, 91) > base
    assert L.sqz_baf_grid_workspace_bytes(0, 0, 0, 0) > 0
    # preferred = minimum + scratch: 8 B x 21 x N per column for the split kernel ...
    b2 = L.sqz_baf_grid_workspace_bytes(3000, 1, 61, 91)
    pref = L.sqz_baf_grid_workspace_preferred_bytes(3000, 1, 61, 91, 91, 21, 0)
    assert pref - b2 == pytest.approx(91 * 21 * 3000 * 8, abs=512)
    # ... 8 B x 8 x N per column in whole 32-column groups for the column-per-lane kernel
    b3 = L.sqz_baf_grid_workspace_bytes(100000, 1, 601, 901)
    p3 = L.sqz_baf_grid_workspace_preferred_bytes(100000, 1, 601, 901, 901, 8, 0)
    assert p3 - b3 == pytest.approx(29 * 32 * 8 * 100000 * 8, rel=1e-6)
    # a column shard of the same grid keeps the kernel of the whole grid
    b8 = L.sqz_baf_grid_workspace_bytes(100000, 1, 601, 12)
    p8 = L.sqz_baf_grid_workspace_preferred_bytes(100000, 1, 601, 901, 12, 8, 0)
    assert p8 - b8 == pytest.approx(32 * 8 * 100000 * 8, rel=1e-6)
    # the cap bounds the scratch, never the minimum
    capped = L.sqz_baf_grid_workspace_preferred_bytes(100000, 1, 601, 901, 901, 8, 1 << 30)
    assert b3 < capped <= b3 + (1 << 30) + 256
    # out-of-range copy-number counts fall back to the minimum
    assert L.sqz_baf_grid_workspace_preferred_bytes(30000, 1, 61, 91, 91, 0, 0) == base
    # a unit shard: the groups it touches (69 units of 19 tiles: at most 5 groups)
    pl = np.linspace(1.0, 7.0, 601)
    u = L.sqz_baf_grid_units_workspace_bytes(100000, 1, _lib.ptr(pl), 601, 901, 0, 7, 2, 0,
                                             69, 138, 0)
    assert 4 * 32 * 8 * 100000 * 8 <= u <= 6 * 32 * 8 * 100000 * 8 + (1 << 30)
    assert L.sqz_baf_grid_units_workspace_bytes(100000, 1, _lib.ptr(pl), 601, 901, 0, 7, 2, 0,
                                                10, 5, 0) == 0  # bad range
