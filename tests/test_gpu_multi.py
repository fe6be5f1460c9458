"""One process driving several GPUs through the host entry points (`n_gpus` > 1): a single
sample is cut into units of the grid (results stored into device 0's buffer over NVLink),
a cohort by sample.  Needs two visible devices; the one-GPU box skips these."""
import numpy as np
import pytest

import sequenza_b200 as S
from sequenza_b200 import _lib
from sequenza_b200.synth import baf_args, synth_segments

pytestmark = pytest.mark.gpu
PRI = {"CN": [2], "value": [2]}


def _need_two():
    if _lib.lib().sqz_device_count() < 2:
        pytest.skip("needs two GPUs")


@pytest.mark.parametrize("n_seg,grid,cn_max", [(4000, (0.01, 0.1), 7), (4000, (0.01, 0.1), 20),
                                               (20000, (0.002, 0.02), 7)])
def test_single_sample_unit_shards_are_bit_identical(n_seg, grid, cn_max):
    """The bits of a grid point do not depend on the number of GPUs (the kernel, its tiles
    and its segment chunks are planned from the whole grid)."""
    _need_two()
    seg = synth_segments(n_seg, seed=71)
    ce, pl = S.r_seq(0.1, 1, grid[0]), S.r_seq(1, 7, grid[1])
    kw = dict(cellularity=ce, ploidy=pl, priors_table=PRI, CNt_max=cn_max, **baf_args(seg))
    one = S.baf_grid_sums(n_gpus=1, **kw)
    for n in sorted({2, _lib.lib().sqz_device_count()}):
        for _ in range(2):  # the second call reuses the cached buffers
            many = S.baf_grid_sums(n_gpus=n, **kw)
            np.testing.assert_array_equal(many, one)


@pytest.mark.parametrize("sizes", [[300, 0, 1200, 50, 700], [1] * 9, [0, 0, 40], [2500] * 16])
def test_cohort_sample_shards(sizes):
    _need_two()
    ce, pl = S.r_seq(0.1, 1, 0.05), S.r_seq(1, 5, 0.25)
    samples = []
    for k, n in enumerate(sizes):
        a = baf_args(synth_segments(n, seed=80 + k, cellularity=0.3 + 0.05 * (k % 10)))
        samples.append(a)
    one = S.baf_model_fit_batch(samples, ce, pl, n_gpus=1, priors_table=PRI)
    for n in (2, _lib.lib().sqz_device_count()):
        many = S.baf_model_fit_batch(samples, ce, pl, n_gpus=n, priors_table=PRI)
        assert len(many) == len(one)
        for a, b in zip(many, one):
            np.testing.assert_allclose(a["lpp"], b["lpp"], rtol=1e-12, atol=0)
