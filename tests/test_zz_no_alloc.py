"""sqz_baf_grid_dev_noalloc: the grid fit for callers that are not allowed to allocate
device memory (a latency-bounded loop where cudaMalloc's driver lock and implicit
synchronisation are unacceptable, a plugin whose host application forbids allocations
after start-up, a process that has given all free HBM to another pool).  It runs the
same kernels as sqz_baf_grid_dev with the scratch in a fixed 16 MiB array inside the
library instead of a caller workspace.

Collected last on purpose (file name): if it fails, `pytest -x` has already run every
other test.  It must pass alone in a fresh process and after the rest of the suite."""
import numpy as np
import pytest

import sequenza_b200 as S
from sequenza_b200 import _lib
from sequenza_b200.synth import baf_args, synth_segments

pytestmark = pytest.mark.gpu
RTOL = 1e-9  # the LPP tolerance of the other GPU parity tests
PRI = {"CN": [2], "value": [2]}


def test_noalloc_matches_workspace_entry_point_twice_in_a_row(oracle):
    """Two consecutive calls on different inputs and sizes (the second is smaller, so
    anything the first call left in the shared scratch would show up) against the
    existing workspace-taking entry point, through the C ABI -- and, as the third leg,
    against the CPU oracle on a window of grid points around the fitted solution."""
    import torch
    from sequenza_b200.dist import BafGridDevice, run_noalloc
    ce, pl = S.r_seq(0.1, 1, 0.01), S.r_seq(1, 7, 0.1)
    cases = [(synth_segments(6000, seed=61), dict(CNt_max=20, priors_table=PRI)),
             (synth_segments(450, cellularity=0.4, ploidy=3.3, seed=62),
              dict(CNt_max=7, priors_table=PRI, ratio_priority=True))]
    for seg, kw in cases:
        dev = BafGridDevice(**baf_args(seg))
        want = dev.run(pl, ce, **kw)                    # sqz_baf_grid_dev (workspace)
        out = torch.full_like(want, float("nan"))       # caller-owned output buffer
        n_alloc = torch.cuda.memory_stats()["allocation.all.allocated"]
        got = run_noalloc(dev, pl, ce, out=out, **kw)   # sqz_baf_grid_dev_noalloc
        torch.cuda.synchronize()
        # the wrapper made no torch allocation either (cumulative allocation count)
        assert torch.cuda.memory_stats()["allocation.all.allocated"] == n_alloc
        a, b = got[0].cpu().numpy(), want[0].cpu().numpy()
        assert np.isfinite(b).all()
        np.testing.assert_allclose(a, b, rtol=RTOL, atol=0)
        assert int(np.argmax(a)) == int(np.argmax(b))
        k = int(np.argmax(a))
        g0 = max(0, min(k - 32, a.size - 64))
        ref = oracle.baf_grid_sums(cellularity=ce, ploidy=pl, g0=g0, g1=g0 + 64,
                                   **baf_args(seg), **kw)
        np.testing.assert_allclose(a[g0:g0 + 64], ref, rtol=RTOL, atol=0)
        assert g0 + int(np.argmax(ref)) == k
    # a column shard, and a degenerate input that goes through the repair kernel
    seg = synth_segments(300, seed=63)
    seg["sd_ratio"][5] = 0.0
    dev = BafGridDevice(**baf_args(seg))
    want = dev.run(pl, ce, 40, 60, priors_table=PRI)[0].cpu().numpy()
    got = run_noalloc(dev, pl, ce, 40, 60, priors_table=PRI)[0].cpu().numpy()
    assert np.isneginf(want).all() and np.array_equal(got, want)


def test_noalloc_refuses_problems_beyond_its_scratch():
    import torch
    from sequenza_b200.dist import BafGridDevice, run_noalloc
    ce, pl = S.r_seq(0.1, 1, 0.001), S.r_seq(1, 7, 0.01)  # 901 x 601 grid
    need = _lib.lib().sqz_baf_grid_workspace_bytes(2000, 1, pl.size, ce.size)
    assert need > 16 << 20
    dev = BafGridDevice(**baf_args(synth_segments(2000, seed=64)))
    out = torch.zeros((1, ce.size * pl.size), dtype=torch.float64, device=dev.device)
    with pytest.raises(RuntimeError, match="16 MiB"):
        run_noalloc(dev, pl, ce, out=out)
    assert float(out.abs().sum()) == 0.0  # refused before touching the output
