"""Device-resident entry point (sqz_baf_grid_dev) under torch plumbing: column shards
computed separately and concatenated equal the one-call result."""
import numpy as np
import pytest

import sequenza_b200 as S
from sequenza_b200.synth import baf_args, synth_segments

pytestmark = pytest.mark.gpu
PRI = {"CN": [2], "value": [2]}


def test_dev_shards_equal_host_call():
    import torch
    from sequenza_b200.dist import BafGridDevice, gather_columns, shard_columns
    seg = synth_segments(5000, seed=31)
    ce, pl = S.r_seq(0.1, 1, 0.01), S.r_seq(1, 7, 0.1)
    host = S.baf_grid_sums(cellularity=ce, ploidy=pl, priors_table=PRI, **baf_args(seg))
    dev = BafGridDevice(**baf_args(seg))
    full = dev.run(pl, ce, priors_table=PRI)
    torch.cuda.synchronize()
    np.testing.assert_array_equal(full[0].cpu().numpy(), host)
    parts = []
    for r in range(4):
        c0, c1 = shard_columns(ce.size, 4, r)
        parts.append(dev.run(pl, ce, c0, c1, priors_table=PRI)[0].cpu().numpy())
    np.testing.assert_array_equal(np.concatenate(parts), host)
    one = gather_columns(full, ce.size, pl.size, 1, 0)
    np.testing.assert_array_equal(one[0].cpu().numpy(), host)


def test_distributed_fit_single_rank():
    from sequenza_b200.dist import baf_model_fit_distributed
    seg = synth_segments(700, seed=32)
    ce, pl = S.r_seq(0.1, 1, 0.05), S.r_seq(1, 7, 0.5)
    a = baf_model_fit_distributed(ce, pl, priors_table=PRI, **baf_args(seg))
    b = S.baf_model_fit(cellularity=ce, ploidy=pl, priors_table=PRI, **baf_args(seg))
    np.testing.assert_array_equal(a["lpp"], b["lpp"])


def test_unit_shards_equal_host_call():
    """sqz_baf_grid_dev_units: every world size's unit ranges, written into one whole-grid
    buffer, reproduce the one-call result bit for bit (split, pruned and column-per-lane
    kernels)."""
    import torch
    from sequenza_b200.dist import BafGridDevice, grid_plan, shard_units
    seg = synth_segments(5000, seed=33)
    for ce, pl, cn in ((S.r_seq(0.1, 1, 0.01), S.r_seq(1, 7, 0.1), 20),     # split
                       (S.r_seq(0.1, 1, 0.01), S.r_seq(1, 7, 0.1), 7),      # pruned
                       (S.r_seq(0.1, 1, 0.004), S.r_seq(1, 7, 0.02), 7)):   # column per lane
        kw = dict(priors_table=PRI, CNt_max=cn)
        host = S.baf_grid_sums(cellularity=ce, ploidy=pl, **kw, **baf_args(seg))
        dev = BafGridDevice(**baf_args(seg))
        plan = grid_plan(5000, 1, pl, ce.size, CNt_max=cn)
        for world in (1, 3, 8):
            buf = torch.full((1, ce.size * pl.size), float("nan"), dtype=torch.float64,
                             device=dev.device)
            for r in range(world):
                u0, u1 = shard_units(plan["units"], world, r)
                dev.run_units(pl, ce, u0, u1, out=buf, **kw)
            np.testing.assert_array_equal(buf[0].cpu().numpy(), host)
        # an empty range is a no-op, a range beyond the grid an error
        dev.run_units(pl, ce, 5, 5, out=buf, **kw)
        with pytest.raises(ValueError):
            dev.run_units(pl, ce, 0, plan["units"] + 1, out=buf, **kw)
