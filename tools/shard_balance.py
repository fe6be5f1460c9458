"""Config 3 on ONE GPU, shard by shard: how long each rank's unit range of an N-way split
takes, and how the cost is spread over the 29 column groups (19 units each).
python tools/shard_balance.py [world ...]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import sequenza_b200 as S  # noqa: E402
from sequenza_b200.dist import BafGridDevice, grid_plan, shard_units  # noqa: E402
from sequenza_b200.synth import baf_args, synth_segments  # noqa: E402

worlds = [int(a) for a in sys.argv[1:]] or [8]
seg = synth_segments(100000)
ce, pl = S.r_seq(0.1, 1, 0.001), S.r_seq(1, 7, 0.01)
dev = BafGridDevice(**baf_args(seg))
pri = {"CN": [2], "value": [2]}
plan = grid_plan(100000, 1, pl, ce.size)
buf = torch.zeros((1, ce.size * pl.size), dtype=torch.float64, device=dev.device)


def timed(u0, u1, reps=5):
    for _ in range(2):
        dev.run_units(pl, ce, u0, u1, out=buf, priors_table=pri)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        dev.run_units(pl, ce, u0, u1, out=buf, priors_table=pri)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


whole = timed(0, plan["units"])
print("whole grid: %.3f ms (%d units, %d chunks)" % (whole, plan["units"], plan["chunks"]))
for world in worlds:
    t = [timed(*shard_units(plan["units"], world, r)) for r in range(world)]
    print("world %d: per rank ms %s; max %.3f, mean %.3f, ideal (whole / world) %.3f -> "
          "efficiency of the slowest %.3f" % (world, ["%.3f" % x for x in t], max(t),
                                              sum(t) / world, whole / world,
                                              whole / world / max(t)))
if os.environ.get("SKIP_GROUPS"):
    sys.exit(0)
nPT = plan["nPT"]
g = [timed(k * nPT, (k + 1) * nPT, reps=3) for k in range(plan["units"] // nPT)]
print("per column group (19 units): " + " ".join("%.2f" % x for x in g))
