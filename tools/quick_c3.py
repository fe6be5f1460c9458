"""Config 3 (1e5 segments, 901 x 601, CNt 0..7) on one GPU, kernel phases by CUDA events:
python tools/quick_c3.py [steps]"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import sequenza_b200 as S  # noqa: E402
from sequenza_b200.dist import BafGridDevice  # noqa: E402
from sequenza_b200.synth import baf_args, synth_segments  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
scale = float(os.environ.get("SD_SCALE", "1"))
C2 = os.environ.get("WORKLOAD", "c3") == "c2"  # BASELINE config 2 instead of config 3
seg = synth_segments(30000 if C2 else 100000)
seg["sd_ratio"] = seg["sd_ratio"] * scale
seg["sd_Bf"] = seg["sd_Bf"] * scale
ce, pl = (S.r_seq(0.1, 1, 0.01), S.r_seq(1, 7, 0.1)) if C2 else \
    (S.r_seq(0.1, 1, 0.001), S.r_seq(1, 7, 0.01))
dev = BafGridDevice(**baf_args(seg))
L = S.lib()
pri = {"CN": [2], "value": [2]}
if C2:
    import functools
    dev.run = functools.partial(dev.run, CNt_max=20)
for _ in range(3):
    out = dev.run(pl, ce, priors_table=pri)
torch.cuda.synchronize()
L.sqz_profile_enable(1)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(steps):
    out = dev.run(pl, ce, priors_table=pri)
e1.record()
torch.cuda.synchronize()
ms3, nc = (C.c_double * 3)(), C.c_int(0)
path, ne, nl = C.c_int(0), C.c_ulonglong(0), C.c_longlong(0)
L.sqz_profile_counters(C.byref(path), C.byref(ne), C.byref(nl))
L.sqz_profile_collect(ms3, C.byref(nc))
k = int(torch.argmax(out))
print(("config2" if C2 else "config3") + " sd x%g: %.3f ms per fit (prep %.3f, pre-pass + grid kernel %.3f, finalize %.3f), "
      "path %d, candidates per (segment, tile) %.3f, argmax %d, sum %.17g, env %s" % (
          scale, e0.elapsed_time(e1) / steps, ms3[0] / nc.value, ms3[1] / nc.value,
          ms3[2] / nc.value, path.value, ne.value / (nc.value * float(len(seg["Bf"])) * ce.size * max(1, -(-pl.size // max(1, int(os.environ.get("SQZ_PT", "32")))))) , k,
          float(out.sum()), {k: v for k, v in os.environ.items() if k.startswith("SQZ_")}))
