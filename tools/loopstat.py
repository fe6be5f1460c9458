#!/usr/bin/env python3
"""Static instruction mix of the hottest loops of baf_grid_fast_kernel<PT,false> (from cuobjdump)."""
import collections, re, subprocess, sys
import os
KERNEL = os.environ.get("KERNEL", "baf_grid_split_kernel")
so = "sequenza_b200/libsequenza_b200.so"
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
for pt in sys.argv[1:] or ["16", "32"]:
    on, ins = False, []
    for l in txt.splitlines():
        if "Function :" in l:
            on = ("%sILi%sELb0E" % (KERNEL, pt)) in l
        if on:
            m = re.match(r"\s+/\*([0-9a-f]+)\*/\s+(.*?);", l)
            if m:
                ins.append((int(m.group(1), 16), m.group(2)))
    loops = []
    for a, t in ins:
        m = re.search(r"BRA\s+(0x[0-9a-f]+)", t)
        if m and int(m.group(1), 16) < a:
            tgt = int(m.group(1), 16)
            body = [x for x in ins if tgt <= x[0] <= a]
            loops.append((sum("DFMA" in b[1] for b in body), tgt, a, body))
    for nd, tgt, a, body in sorted(loops, key=lambda x: -x[0])[:3]:
        c = collections.Counter()
        for _, t in body:
            op = t.split()[1] if t.startswith("@") else t.split()[0]
            c[op.split(".")[0]] += 1
        print("PT", pt, "loop", hex(tgt), hex(a), "instrs", len(body), dict(c.most_common(12)))
