#!/usr/bin/env python3
"""gpurun_out/ ncu artefacts of round 2 -> tracked files under profiles/:
   r02_launches_config{2,3}.csv   ncu launch lists (copied)
   r02_flops_config{2,3}.csv      executed fp64 instruction counters per kernel (copied)
   r02_traffic.json               DRAM bytes per launch of the dominant kernels (bench.py reads it)
   r02_ncu_summary.md             launch-list shares, full-capture metrics, stall hot spots,
                                  the reconciliation of bench.py's instruction model with the
                                  counters
python tools/profiles_r02.py <job prefix, e.g. gpurun_out/j7> [git commit of the capture]"""
import collections
import csv
import json
import os
import shutil
import subprocess
import sys

pre = sys.argv[1]
commit = sys.argv[2] if len(sys.argv) > 2 else subprocess.run(
    ["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
out = ["# ncu evidence, round 2 (captured at commit %s)\n" % commit]
EVALS = {"c3": 901 * 601 * 1e5, "c2": 91 * 61 * 3e4}
NAME = {"c3": "config 3 (1e5 segments, CNt 0..7, 901 x 601)",
        "c2": "config 2 (3e4 segments, CNt 0..20, 91 x 61)"}


def launches(tag):
    path = "%s_launches_%s.csv" % (pre, tag)
    if not os.path.exists(path):
        return
    shutil.copy(path, "profiles/r02_launches_config%s.csv" % tag[1])
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    hdr = rows[0]
    ik, iv, ig, ib = (hdr.index(c) for c in ("Kernel Name", "Metric Value", "Grid Size", "Block Size"))
    agg = collections.OrderedDict()
    for r in rows[1:]:
        k = r[ik].split("(")[0]
        agg.setdefault(k, {"ns": [], "grid": r[ig], "block": r[ib]})["ns"].append(
            float(r[iv].replace(",", "")))
    out.append("## Launch list, %s\n" % NAME[tag])
    out.append("`ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv python "
               "bench.py --steps N --warmup 3%s` after the same command exited 0 without ncu. "
               "Per-launch times are cold-cache and serialised: compare shares.\n"
               % (" --workload config2" if tag == "c2" else ""))
    out.append("| kernel | launches | grid | block | avg us |")
    out.append("|---|---|---|---|---|")
    for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1]["ns"])):
        out.append("| `%s` | %d | %s | %s | %.1f |" % (k, len(v["ns"]), v["grid"], v["block"],
                                                     sum(v["ns"]) / len(v["ns"]) / 1e3))
    fit = {k: sum(v["ns"]) / len(v["ns"]) for k, v in agg.items() if "baf_" in k}
    want = ("baf_prep", "baf_colterm", "baf_grid_colwarp", "baf_grid_split", "baf_grid_finalize",
            "baf_grid_faithful")
    # config 3's line also runs the config 2 and worst-case extras: keep the fit's own kernels
    tot = sum(fit.values())
    out.append("\nShare of one fit (average launch of each `baf_*` kernel): " + ", ".join(
        "`%s` %.1f%%" % (k.replace("void ", "").replace("sqz::", ""), 100 * v / tot)
        for k, v in sorted(fit.items(), key=lambda kv: -kv[1])) + "\n")


def full(tag):
    path = "%s_%s_colwarp.ncu-rep" % (pre, tag)
    if not os.path.exists(path):
        return None
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True,
                         text=True).stdout
    r = list(csv.reader(raw.splitlines()))
    h, u, v = r[0], r[1], r[2]
    m = dict(zip(h, zip(u, v)))
    want = ["Kernel Name", "gpu__time_duration.sum", "launch__grid_size", "launch__block_size",
            "launch__registers_per_thread", "launch__occupancy_limit_registers",
            "launch__occupancy_limit_shared_mem", "sm__cycles_elapsed.avg.per_second",
            "sm__cycles_active.avg", "sm__cycles_elapsed.avg",
            "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
            "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
            "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
            "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
            "sm__warps_active.avg.pct_of_peak_sustained_active", "dram__bytes_read.sum",
            "dram__bytes_write.sum", "lts__t_sector_hit_rate.pct",
            "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
            "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio"]
    out.append("## Full capture of the dominant kernel, %s\n" % NAME[tag])
    out.append("`ncu --set full --clock-control none --import-source on -k regex:baf_grid_colwarp "
               "-s 2 -c 1 python tools/quick_c3.py`%s\n" % (" with WORKLOAD=c2" if tag == "c2" else ""))
    out.append("| metric | unit | value |")
    out.append("|---|---|---|")
    for name in want:
        if name in m:
            out.append("| %s | %s | %s |" % (name, m[name][0], m[name][1]))
    hot = subprocess.run([sys.executable, "tools/hotspots.py", path, "2"], capture_output=True,
                         text=True).stdout
    out.append("\nRegions of the SASS with more than 2 %% of the stall samples (`tools/hotspots.py`: "
               "first address, instructions, executions, share of samples and of issued "
               "instructions, top stall reasons, opcode mix):\n\n```\n%s```\n" % hot)

    def gb(name):
        unit, val = m[name]
        f = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[unit]
        return float(val.replace(",", "")) * f
    return {"kernel": m["Kernel Name"][1].split("(")[0], "kernel_path": 3, "commit": commit,
            "dram_bytes_per_launch": gb("dram__bytes_read.sum") + gb("dram__bytes_write.sum"),
            "dram_bytes_read": gb("dram__bytes_read.sum"),
            "dram_bytes_write": gb("dram__bytes_write.sum"),
            "warp_instructions_per_launch": float(m["smsp__inst_executed.sum"][1].replace(",", "")),
            "issue_active_pct": float(
                m["smsp__issue_active.avg.pct_of_peak_sustained_active"][1].replace(",", "")),
            "source": "ncu --set full --clock-control none, one launch"}


def flops(tag):
    path = "%s_flops_%s.csv" % (pre, tag)
    if not os.path.exists(path):
        return
    shutil.copy(path, "profiles/r02_flops_config%s.csv" % tag[1])
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    hdr = [r for r in rows if "Metric Name" in r][0]
    ik, im, iv = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value")
    k = collections.defaultdict(dict)
    for r in rows:
        if r is hdr:
            continue
        k[r[ik].split("(")[0].replace("void ", "")][r[im]] = float(r[iv].replace(",", ""))
    out.append("## Executed fp64 work from the counters, %s\n" % NAME[tag])
    out.append("`ncu --metrics smsp__sass_thread_inst_executed_op_{dfma,dmul,dadd}_pred_on.sum,"
               "smsp__inst_executed_pipe_fp64.sum,...` on one launch of the pre-pass and of the grid "
               "kernel.\n")
    out.append("| kernel | DFMA | DMUL | DADD | flops (DFMA = 2) | fp64-pipe warp instructions |")
    out.append("|---|---|---|---|---|---|")
    tot_f = tot_i = 0.0
    for name, d in k.items():
        f = 2 * d["smsp__sass_thread_inst_executed_op_dfma_pred_on.sum"] + \
            d["smsp__sass_thread_inst_executed_op_dmul_pred_on.sum"] + \
            d["smsp__sass_thread_inst_executed_op_dadd_pred_on.sum"]
        tot_f += f
        tot_i += d["smsp__inst_executed_pipe_fp64.sum"]
        out.append("| `%s` | %.4g | %.4g | %.4g | %.4g | %.4g |" % (
            name[:40], d["smsp__sass_thread_inst_executed_op_dfma_pred_on.sum"],
            d["smsp__sass_thread_inst_executed_op_dmul_pred_on.sum"],
            d["smsp__sass_thread_inst_executed_op_dadd_pred_on.sum"], f,
            d["smsp__inst_executed_pipe_fp64.sum"]))
    ev = EVALS[tag]
    out.append("\nPer evaluation (%.4g evaluations per fit): **%.2f flops, %.2f fp64-pipe thread "
               "instructions** from the counters.  bench.py's analytic model "
               "(`roofline.flops_per_eval_executed`, `fp64_issue_slots_per_eval`) is checked "
               "against these in DESIGN.md 4.1.\n" % (ev, tot_f / ev, 32 * tot_i / ev))


def other(path, title, cmd, bytes_note):
    """launch list with DRAM bytes of a secondary workload (committed CSV under profiles/)"""
    if not os.path.exists(path):
        return
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    hdr = [r for r in rows if "Metric Name" in r][0]
    ik, im, iv, iid = (hdr.index(c) for c in ("Kernel Name", "Metric Name", "Metric Value", "ID"))
    per = collections.OrderedDict()
    for r in rows:
        if r is hdr:
            continue
        per.setdefault((r[iid], r[ik].split("(")[0]), {})[r[im]] = float(r[iv].replace(",", ""))
    agg = collections.OrderedDict()
    for (_, k), d in per.items():
        a = agg.setdefault(k, {"n": 0, "ns": 0.0, "rd": 0.0, "wr": 0.0})
        a["n"] += 1
        a["ns"] += d.get("gpu__time_duration.sum", 0.0)
        a["rd"] += d.get("dram__bytes_read.sum", 0.0)
        a["wr"] += d.get("dram__bytes_write.sum", 0.0)
    out.append("## Launch list, %s\n" % title)
    out.append("`%s` (`%s`). %s\n" % (cmd, path, bytes_note))
    out.append("| kernel | launches | avg us | DRAM read MB | DRAM write MB | DRAM GB/s |")
    out.append("|---|---|---|---|---|---|")
    for k, a in agg.items():
        n = a["n"]
        out.append("| `%s` | %d | %.1f | %.1f | %.1f | %.0f |" % (
            k.replace("void ", "").replace("sqz::", ""), n, a["ns"] / n / 1e3, a["rd"] / n / 1e6,
            a["wr"] / n / 1e6, (a["rd"] + a["wr"]) / a["ns"] if a["ns"] else 0.0))
    out.append("")


traffic = {}
for tag in ("c3", "c2"):
    launches(tag)
    t = full(tag)
    if t:
        traffic["config%s" % tag[1]] = t
    flops(tag)
M3 = "ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum " \
     "--clock-control none"
other("profiles/r02_launches_config4.csv",
      "config 4 (mufreq.bayes, 1e7 positions, scalar weights 100 / 100, T = 230)",
      M3 + " -k regex:mufreq -c 12 python bench.py --workload config4",
      "Algorithmic bytes of the lookup: 16 B in + 32 B out per position = 480 MB (part of the "
      "output is still in L2 when the kernel ends).")
other("profiles/r02_launches_segstats.csv",
      "segment statistics (2.4e7 seqz rows, 2,400 segments, 30 % het)",
      M3 + " -k regex:segstat -c 8 python bench.py --workload segstats",
      "Algorithmic bytes of the reduction: 49 B per row = 1.18 GB; it also writes the compact "
      "copy of the het rows (24 B each) that the density kernel reads back.")
if traffic:
    json.dump(traffic, open("profiles/r02_traffic.json", "w"), indent=1)
open("profiles/r02_ncu_summary.md", "w").write("\n".join(out) + "\n")
print("wrote profiles/r02_ncu_summary.md, r02_traffic.json")
