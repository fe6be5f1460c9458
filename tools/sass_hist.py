#!/usr/bin/env python3
"""cuobjdump -sass of the built library -> profiles/r02_sass_histogram.md: per kernel the
number of SASS instructions and the count of each opcode family that matters for the
roofline argument (fp64 pipe, fp32, integer/select, shared/global memory, TMA bulk copies and
mbarrier operations, warp collectives, local-memory spills).
python tools/sass_hist.py [sequenza_b200/libsequenza_b200.so]"""
import collections
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else "sequenza_b200/libsequenza_b200.so"
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
commit = subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True,
                        text=True).stdout.strip()
FAM = collections.OrderedDict([
    ("fp64 pipe", r"^(DFMA|DMUL|DADD|DSETP|DMNMX)"),
    ("MUFU", r"^MUFU"),
    ("fp32", r"^(FFMA|FMUL|FADD|FMNMX|FSETP|FCHK)"),
    ("FSEL/SEL", r"^(FSEL|SEL)"),
    ("integer", r"^(IMAD|IADD3|VIADD|LOP3|SHF|LEA|ISETP|VIMNMX|IABS|POPC|FLO|BREV|PRMT|I2F|F2I|F2F|I2I|MOV|UMOV|UIADD3|ULEA|ULOP3|USHF|UISETP|UIMAD|USEL|UPRMT|CS2R|S2R|S2UR|R2UR)"),
    ("LDS/STS", r"^(LDS|STS)"),
    ("LDG/STG", r"^(LDG|STG|LD\.|ST\.|LDC|LDCU|ATOM|ATOMG|ATOMS|RED)"),
    ("LDL/STL (spill)", r"^(LDL|STL)"),
    ("UBLKCP (TMA bulk)", r"^UBLKCP"),
    ("SYNCS (mbarrier)", r"^SYNCS"),
    ("LDGSTS (cp.async)", r"^LDGSTS"),
    ("warp collective", r"^(SHFL|VOTE|REDUX|CREDUX|MATCH|WARPSYNC)"),
    ("BAR", r"^BAR"),
    ("branch", r"^(BRA|BSSY|BSYNC|EXIT|RET|CALL|BRX|JMP)"),
])
kern = collections.OrderedDict()
cur = None
for line in txt.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = m.group(1)
        kern[cur] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\w+\s+)?([A-Z0-9_.]+)", line)
    if m and cur:
        op = m.group(1)
        kern[cur]["total"] += 1
        for name, pat in FAM.items():
            if re.match(pat, op):
                kern[cur][name] += 1
                break
        else:
            kern[cur]["other"] += 1


def demangle(names):
    r = subprocess.run(["cu++filt"] + names, capture_output=True, text=True)
    return r.stdout.splitlines() if r.returncode == 0 else names


names = list(kern)
pretty = demangle(names)
out = ["# SASS instruction histogram per kernel (sm_100a, commit %s)\n" % commit,
       "`cuobjdump -sass %s`, static counts (every instantiation that is compiled in); "
       "`tools/sass_hist.py` regenerates this file.\n" % lib,
       "UBLKCP = `cp.async.bulk` (1-D TMA bulk copy global -> shared, no multicast), SYNCS = "
       "`mbarrier` arrive / expect_tx / try_wait.\n",
       "| kernel | total | " + " | ".join(FAM) + " | other |",
       "|---|---|" + "---|" * (len(FAM) + 1)]
for n, p in sorted(zip(names, pretty), key=lambda t: t[1]):
    k = kern[n]
    short = re.sub(r"\(.*", "", p).replace("void ", "").replace("sqz::", "")
    short = re.sub(r"\(anonymous namespace\)::", "", short)
    out.append("| `%s` | %d | " % (short, k["total"]) +
               " | ".join(str(k[f]) if k[f] else "" for f in FAM) + " | %d |" % k["other"])
open("profiles/r02_sass_histogram.md", "w").write("\n".join(out) + "\n")
print("\n".join(out[:12]))
