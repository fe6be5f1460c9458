import sys, time; sys.path.insert(0, ".")
import numpy as np, sequenza_b200 as S
from sequenza_b200.synth import synth_segments, baf_args
PRI = {"CN": [2], "value": [2]}
# very long segment table, small grid: additivity over a split + fast == faithful on a subgrid
seg = synth_segments(3_000_000, seed=9)
ce, pl = np.array([0.3, 0.7, 1.0]), S.r_seq(1.5, 3.5, 0.5)
t = time.perf_counter(); full = S.baf_grid_sums(cellularity=ce, ploidy=pl, priors_table=PRI, **baf_args(seg)); dt = time.perf_counter() - t
h = {k: (v[:1234567] if isinstance(v, np.ndarray) else v) for k, v in baf_args(seg).items()}
r = {k: (v[1234567:] if isinstance(v, np.ndarray) else v) for k, v in baf_args(seg).items()}
parts = S.baf_grid_sums(cellularity=ce, ploidy=pl, priors_table=PRI, **h) + S.baf_grid_sums(cellularity=ce, ploidy=pl, priors_table=PRI, **r)
print("N=3e6: %.1f ms, additivity rel err %.2e" % (dt * 1e3, np.max(np.abs(parts - full) / np.abs(full))))
sub = {k: (v[:200000] if isinstance(v, np.ndarray) else v) for k, v in baf_args(seg).items()}
a = S.baf_grid_sums(cellularity=ce, ploidy=pl, priors_table=PRI, **sub)
b = S.baf_grid_sums(cellularity=ce, ploidy=pl, priors_table=PRI, mode=S.SQZ_MODE_FAITHFUL, **sub)
print("N=2e5 fast vs faithful rel err %.2e (sum magnitude %.3e)" % (np.max(np.abs(a - b) / np.abs(b)), np.abs(b).max()))
# many tiny samples in one launch
samples = [baf_args(synth_segments(int(n), seed=int(n))) for n in np.random.default_rng(0).integers(1, 60, 3000)]
t = time.perf_counter(); fits = S.baf_model_fit_batch(samples, cellularity=S.r_seq(0.1, 1, 0.05), ploidy=S.r_seq(1, 7, 0.25), priors_table=PRI); dt = time.perf_counter() - t
print("3000 tiny samples in one call: %.1f ms, all normalised: %s" % (dt * 1e3, all(abs(f["lpp"].sum() - 1) < 1e-9 for f in fits)))
# repeated calls are bit-reproducible
x = [S.baf_grid_sums(cellularity=ce, ploidy=pl, priors_table=PRI, **sub) for _ in range(3)]
print("repeatable:", all(np.array_equal(x[0], y) for y in x[1:]))
