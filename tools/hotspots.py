#!/usr/bin/env python3
"""Region-level summary of an ncu source page: python tools/hotspots.py <rep> [min_pct]"""
import csv, math, subprocess, sys
rep = sys.argv[1]; minpct = float(sys.argv[2]) if len(sys.argv) > 2 else 1.5
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr = rows[1]; data = rows[2:]
ia, isrc, isamp, iex = (hdr.index(c) for c in ("Address", "Source", "# Samples", "Instructions Executed"))
cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
idx = {c: hdr.index(c) for c in cols}
tot = sum(int(r[isamp]) for r in data); totex = sum(int(r[iex]) for r in data)
print("kernel:", rows[0][1][:90]); print("samples", tot, "warp-instr", totex)
groups = []; cur = None
for i, r in enumerate(data):
    ex = int(r[iex]); key = round(math.log10(ex + 1), 1)
    if cur is None or abs(key - cur) > 0.12:
        cur = key; groups.append([i, 0, 0, ex, 0])
    groups[-1][1] += 1; groups[-1][2] += int(r[isamp]); groups[-1][4] += ex
for g in groups:
    if g[2] > tot * minpct / 100:
        st = {c: 0 for c in cols}; ops = {}
        for r in data[g[0]:g[0] + g[1]]:
            for c in cols: st[c] += int(r[idx[c]] or 0)
            o = r[isrc].split()[1] if r[isrc].startswith("@") else r[isrc].split()[0]
            ops[o.split(".")[0]] = ops.get(o.split(".")[0], 0) + 1
        top = sorted(st.items(), key=lambda kv: -kv[1])[:4]
        topo = sorted(ops.items(), key=lambda kv: -kv[1])[:6]
        print("%s n=%4d exec=%-10d samples %5.1f%% instr %5.1f%%  %s  %s" % (
            data[g[0]][ia][-5:], g[1], g[3], 100 * g[2] / tot, 100 * g[4] / totex,
            [(k[6:], v) for k, v in top], topo))
