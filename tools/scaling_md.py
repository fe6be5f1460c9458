#!/usr/bin/env python3
"""gpurun_out/scale_<workload>_n<N>.json (from tools/scale_run.sh) -> profiles/<tag>_scaling.md
   python tools/scaling_md.py <tag> "<note appended verbatim>" """
import json, sys
tag = sys.argv[1]
note = sys.argv[2] if len(sys.argv) > 2 else ""
out = ["# Scaling on one 8 x B200 box (%s)\n" % tag,
       "`bash tools/scale_run.sh 8` under `gpurun --gpus 8`: `torchrun`, one rank per GPU, NCCL "
       "all-gather of result$LPP; `value` from CUDA events (max over ranks), `e2e` with pinned "
       "host inputs/outputs inside the timing.\n",
       "| workload | GPUs | scaling | ms/step | evals/s (value) | e2e ms/step | e2e evals/s | "
       "efficiency vs N=1 | SM MHz | reasons |", "|---|---|---|---|---|---|---|---|---|---|"]
for w in ("config2", "config3"):
    base = None
    for n in (1, 2, 4, 8):
        try:
            d = json.loads(open("gpurun_out/scale_%s_n%d.json" % (w, n)).read().strip().splitlines()[-1])
        except Exception:
            continue
        base = base or d["value"]
        out.append("| %s | %d | %s | %.3f | %.3e | %.3f | %.3e | %.1f%% | %s | %s |" % (
            w, n, d["scaling"], d["ms_per_step"], d["value"], d["e2e"]["ms_per_step"],
            d["e2e"]["value"], 100 * d["value"] / (base * n), d["clocks"].get("sm_mhz"),
            ",".join(d["clocks"].get("reasons", [])) or "-"))
out.append("\nconfig2 = BASELINE configs[1] weakly scaled (91 cellularity columns x 61 ploidies x 3e4 "
           "segments per GPU, CNt.max = 20); config3 = BASELINE configs[2] (901 x 601 grid, 1e5 "
           "segments, CNt.max = 7) split by cellularity column.\n")
out.append(note + "\n")
open("profiles/%s_scaling.md" % tag, "w").write("\n".join(out))
print("\n".join(out))
