#!/usr/bin/env python3
"""Turns gpurun_out/ ncu artefacts into the tracked summaries under profiles/.

    python tools/summarize_ncu.py <round-tag> <launches.csv>... <full.ncu-rep>... [<bench.json>...]
"""
import collections
import csv
import json
import subprocess
import sys

tag = sys.argv[1]
lists = [a for a in sys.argv[2:] if a.endswith(".csv")]
reps = [a for a in sys.argv[2:] if a.endswith(".ncu-rep")]
bench = [a for a in sys.argv[2:] if a.endswith(".json")]
out = []
for launches in lists:
    rows = [r for r in csv.reader(open(launches)) if len(r) > 5]
    hdr = rows[0]
    ik, iv, ig, ib = (hdr.index(c) for c in ("Kernel Name", "Metric Value", "Grid Size", "Block Size"))
    agg = collections.OrderedDict()
    for r in rows[1:]:
        k = r[ik].split("(")[0]
        agg.setdefault(k, {"ns": [], "grid": r[ig], "block": r[ib]})["ns"].append(
            float(r[iv].replace(",", "")))
    tot = sum(sum(v["ns"]) for v in agg.values())
    c3 = "c3" in launches.split("/")[-1] or "config3" in launches
    out.append("# ncu launch list (%s, %s)\n" % (tag, "config 3" if c3 else "config 2, the default workload"))
    out.append("Command: `ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv "
               "python bench.py %s` (after the same command exited 0 without "
               "ncu).  Times are cold-cache and serialised: compare shares, not absolutes.\n"
               % ("--steps 5 --warmup 3 --workload config3" if c3 else "--steps 20 --warmup 3"))
    out.append("| kernel | launches | grid | block | total ms | avg us | share |")
    out.append("|---|---|---|---|---|---|---|")
    for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1]["ns"])):
        s_ = sum(v["ns"])
        out.append("| `%s` | %d | %s | %s | %.3f | %.1f | %.1f%% |" % (
            k, len(v["ns"]), v["grid"], v["block"], s_ / 1e6, s_ / len(v["ns"]) / 1e3, 100 * s_ / tot))
    step = [k for k in agg if "baf_" in k]
    stot = sum(sum(agg[k]["ns"]) for k in step)
    out.append("\nShare of the fit step (the `baf_*` kernels only):\n")
    for k in step:
        out.append("- `%s`: %.2f%%" % (k, 100 * sum(agg[k]["ns"]) / stot))
    out.append("")

for rep_ in reps:
    raw = subprocess.run(["ncu", "-i", rep_, "--page", "raw", "--csv"], capture_output=True,
                         text=True).stdout
    r = list(csv.reader(raw.splitlines()))
    h, u, v = r[0], r[1], r[2]
    want = ["Kernel Name", "gpu__time_duration.sum", "launch__grid_size", "launch__block_size",
            "launch__registers_per_thread", "launch__occupancy_limit_registers",
            "sm__cycles_elapsed.avg.per_second", "sm__cycles_active.avg", "sm__cycles_elapsed.avg",
            "sm__throughput.avg.pct_of_peak_sustained_elapsed",
            "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
            "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed",
            "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
            "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
            "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
            "smsp__issue_active.avg.pct_of_peak_sustained_active",
            "smsp__inst_executed.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
            "dram__bytes_read.sum", "dram__bytes_write.sum",
            "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
            "smsp__sass_thread_inst_executed_op_dfma_pred_on.sum",
            "smsp__sass_thread_inst_executed_op_dmul_pred_on.sum",
            "smsp__sass_thread_inst_executed_op_dadd_pred_on.sum",
            "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio"]
    out.append("\n# ncu --set full (%s): %s\n" % (tag, rep_.split("/")[-1]))
    out.append("Command: `ncu --set full --clock-control none --import-source on -k "
               "regex:<kernel> -s 3 -c 1 python bench.py [--workload config3] ...` (after the "
               "same command exited 0 without ncu)\n")
    out.append("| metric | unit | value |")
    out.append("|---|---|---|")
    for name in want:
        for hh, uu, vv in zip(h, u, v):
            if hh == name:
                out.append("| %s | %s | %s |" % (hh, uu, vv))
for b in bench:
    for line in open(b):
        if line.startswith("{"):
            d = json.loads(line)
            out.append("\n# bench line (%s, not under ncu): %s\n" % (tag, b.split("/")[-1]))
            out.append("```json\n" + json.dumps(d, indent=1) + "\n```")
open("profiles/%s_summary.md" % tag, "w").write("\n".join(out) + "\n")
print("wrote profiles/%s_summary.md" % tag)
