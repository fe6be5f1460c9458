"""Launch planning and workspace sizing (host code inside the library; no GPU needed):
which kernel path and geometry `sqz_baf_grid_plan` reports for the BASELINE shapes, and
what `sqz_baf_grid_workspace*_bytes` ask for.  CPU only."""
import ctypes as C

import pytest

from sequenza_b200 import _lib


def plan(n, s, nP, nC, cn_max, rp=0):
    out = (C.c_int * 5)()
    _lib.check(_lib.lib().sqz_baf_grid_plan(n, s, nP, nC, 0, cn_max, 2, rp, out))
    return {"PT": out[0], "nPT": out[1], "chunks": out[2], "threads": out[3], "T": out[4]}


@pytest.fixture(autouse=True)
def no_overrides(monkeypatch):
    for k in ("SQZ_PT", "SQZ_THREADS", "SQZ_CHUNKS", "SQZ_SPLIT", "SQZ_PRUNE", "SQZ_COLWARP"):
        monkeypatch.delenv(k, raising=False)


def test_config2_takes_the_split_kernel():
    # 91 x 61 grid, CNt 0..20: a coarse ploidy axis, 128-thread CTAs, 32-wide tiles
    p = plan(30000, 1, 61, 91, 20)
    assert (p["PT"], p["nPT"], p["threads"], p["T"]) == (32, 2, 128, 121)
    assert 1 <= p["chunks"] <= 64


def test_config3_takes_the_column_per_lane_kernel():
    # one warp per CTA, 32-wide tiles, chunk count from the segment table alone
    p = plan(100000, 1, 601, 901, 7)
    assert (p["PT"], p["nPT"], p["threads"], p["chunks"], p["T"]) == (32, 19, 32, 64, 20)
    # an eighth of the columns: four waves of one-warp units with long chunks keep the
    # 32-wide tiles; the same grid over 3,000 segments (65 per chunk) takes 16-wide ones
    q = plan(100000, 1, 601, 113, 7)
    assert (q["PT"], q["threads"], q["chunks"]) == (32, 32, 64)
    assert plan(3000, 1, 601, 181, 7)["PT"] == 16 and plan(3000, 1, 601, 181, 7)["threads"] == 32
    # a shard too narrow to fill the lanes of a 32-column group stays segment-per-lane
    assert plan(100000, 1, 601, 12, 7)["threads"] == 128
    # ratio.priority has no column-per-lane instantiation
    assert plan(100000, 1, 601, 901, 7, rp=1)["threads"] == 128


def test_cohort_and_small_fits():
    # cohort: the samples fill the machine, few chunks (each costs a slab of partials)
    p = plan(512 * 3000, 512, 61, 91, 7)
    assert p["threads"] == 32 and p["PT"] == 16 and p["chunks"] == 4
    # the bundled example's size: segment-per-lane kernel
    assert plan(150, 1, 61, 91, 7)["threads"] in (64, 128)
    # CNt 0..20 on a fine ploidy axis: column-per-lane, 16-wide tiles
    p = plan(30000, 1, 601, 181, 20)
    assert (p["PT"], p["threads"]) == (16, 32)


def test_column_per_lane_chunks_do_not_depend_on_the_shard():
    a = plan(100000, 1, 601, 901, 7)["chunks"]
    for nC in (450, 225, 113):
        assert plan(100000, 1, 601, nC, 7)["chunks"] == a


def test_workspace_sizes():
    L = _lib.lib()
    # minimum: grows with the grid, the segment count and the sample count
    base = L.sqz_baf_grid_workspace_bytes(30000, 1, 61, 91)
    assert base % 256 == 0 and base <= 16 << 20  # fits the no-alloc entry point's scratch
    assert L.sqz_baf_grid_workspace_bytes(30000, 1, 601, 901) > base
    assert L.sqz_baf_grid_workspace_bytes(300000, 1, 61, 91) > base
    assert L.sqz_baf_grid_workspace_bytes(0, 0, 0, 0) > 0
    # preferred = minimum + scratch: 16 B x 21 x N per column for the split kernel ...
    pref = L.sqz_baf_grid_workspace_preferred_bytes(30000, 1, 61, 91, 21, 0)
    assert pref - base == pytest.approx(91 * 21 * 30000 * 16, rel=1e-6)
    # ... 8 B x 8 x N per column in whole 32-column groups for the column-per-lane kernel
    b3 = L.sqz_baf_grid_workspace_bytes(100000, 1, 601, 901)
    p3 = L.sqz_baf_grid_workspace_preferred_bytes(100000, 1, 601, 901, 8, 0)
    assert p3 - b3 == pytest.approx(29 * 32 * 8 * 100000 * 8, rel=1e-6)
    # the cap bounds the scratch, never the minimum
    capped = L.sqz_baf_grid_workspace_preferred_bytes(100000, 1, 601, 901, 8, 1 << 30)
    assert b3 < capped <= b3 + (1 << 30) + 256
    # out-of-range copy-number counts fall back to the minimum
    assert L.sqz_baf_grid_workspace_preferred_bytes(30000, 1, 61, 91, 0, 0) == base
