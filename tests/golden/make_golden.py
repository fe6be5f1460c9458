#!/usr/bin/env python3
"""Regenerates the golden fixtures in this directory from the read-only reference tree.

Run in the build container only (needs /root/reference); the fixtures it writes are
committed so that nothing at test time reads /root/reference:

    python tests/golden/make_golden.py

Sources (paths relative to /root/reference):
  testthat_model.json   tests/testthat/test.model.R:4-55   known answers for
                        theoretical.depth.ratio / .baf / .mufreq and both type matrices
  vignette_rows.json    inst/doc/sequenza.html (table rendered from
                        inst/doc/sequenza.R:58-68): six baf.bayes output rows that
                        sequenza.results produced at the CP.example solution
  cp_example.json       data/CP.example.RData: stored list(ploidy, cellularity, lpp)
                        (man/CP.data.Rd:17-25); R serialisation v2 / XDR decoded below
"""
import gzip
import html
import json
import os
import re
import struct

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


# ---- minimal reader for R's XDR serialisation (RData "RDX2") -------------------------
class RReader:
    def __init__(self, buf):
        self.b, self.i, self.refs = buf, 0, []

    def take(self, n):
        v = self.b[self.i:self.i + n]
        self.i += n
        return v

    def int(self):
        return struct.unpack(">i", self.take(4))[0]

    def item(self):
        flags = self.int()
        ty, has_attr, has_tag = flags & 0xFF, bool(flags & 0x200), bool(flags & 0x400)
        if ty == 254:  # NILVALUE_SXP
            return None
        if ty == 255:  # REFSXP
            return self.refs[(flags >> 8) - 1]
        if ty == 1:  # SYMSXP
            name = self.item()
            self.refs.append(name)
            return name
        if ty == 9:  # CHARSXP
            n = self.int()
            return None if n == -1 else self.take(n).decode("latin1")
        if ty == 2:  # LISTSXP (pairlist): attr, tag, car, cdr
            out = []
            while True:
                attr = self.item() if has_attr else None
                tag = self.item() if has_tag else None
                car = self.item()
                out.append((tag, car))
                flags = self.int()
                ty = flags & 0xFF
                has_attr, has_tag = bool(flags & 0x200), bool(flags & 0x400)
                if ty == 254:
                    return out
                if ty != 2:
                    raise ValueError("unexpected pairlist tail type %d" % ty)
        if ty in (10, 13):  # LGLSXP / INTSXP
            n = self.int()
            val = list(struct.unpack(">%di" % n, self.take(4 * n)))
        elif ty == 14:  # REALSXP
            n = self.int()
            val = list(struct.unpack(">%dd" % n, self.take(8 * n)))
        elif ty == 16:  # STRSXP
            n = self.int()
            val = [self.item() for _ in range(n)]
        elif ty == 19:  # VECSXP
            n = self.int()
            val = [self.item() for _ in range(n)]
        else:
            raise ValueError("unsupported SEXP type %d" % ty)
        attrs = dict(self.item()) if has_attr else {}
        return {"value": val, "attr": attrs}


def read_rdata(path):
    raw = gzip.open(path, "rb").read()
    assert raw[:5] == b"RDX2\n", raw[:5]
    r = RReader(raw)
    r.i = 5
    assert r.take(2) == b"X\n"
    version, writer, min_reader = r.int(), r.int(), r.int()
    top = r.item()
    return {"version": version, "writer": writer, "objects": dict(top)}


def cp_example():
    rd = read_rdata(os.path.join(REF, "data/CP.example.RData"))
    obj = rd["objects"]["CP.example"]
    names = obj["attr"]["names"]["value"]
    parts = dict(zip(names, obj["value"]))
    lpp = parts["lpp"]
    dim = lpp["attr"]["dim"]["value"]
    dn = lpp["attr"]["dimnames"]["value"]
    w = rd["writer"]
    return {
        "source": "data/CP.example.RData (man/CP.data.Rd:17-25)",
        "written_by_R": "%d.%d.%d" % (w >> 16, (w >> 8) & 0xFF, w & 0xFF),
        "names": names,
        "ploidy": parts["ploidy"]["value"],
        "cellularity": parts["cellularity"]["value"],
        "lpp_dim": dim,
        "lpp_rownames": dn[0]["value"],
        "lpp_colnames": dn[1]["value"],
        "lpp_column_major": lpp["value"],
    }


def vignette_rows():
    s = open(os.path.join(REF, "inst/doc/sequenza.html"), encoding="utf8",
             errors="replace").read()
    tables = re.findall(r"<table.*?</table>", s, flags=re.S)
    for t in tables:
        if "depth.ratio" in t and "LPP" in t:
            head = [html.unescape(re.sub(r"<[^>]+>", "", h)).strip()
                    for h in re.findall(r"<th[^>]*>(.*?)</th>", t, flags=re.S)]
            rows = []
            for tr in re.findall(r"<tr[^>]*>(.*?)</tr>", t, flags=re.S):
                cells = [html.unescape(re.sub(r"<[^>]+>", "", c)).strip()
                         for c in re.findall(r"<td[^>]*>(.*?)</td>", tr, flags=re.S)]
                if cells:
                    rows.append(cells)
            return {
                "source": "inst/doc/sequenza.html, kable(head(seg.tab)) from "
                          "inst/doc/sequenza.R:58-68; values rounded to 3 dp by "
                          "inst/doc/sequenza.R:65; produced by sequenza.results "
                          "(R/results.R:128-136: CNt.max = 20, weight.Bf = 1, "
                          "weight.ratio = segment length in Mb, CNn = 2)",
                "columns": head,
                "rows": rows,
            }
    raise RuntimeError("segment table not found in vignette")


def testthat_model():
    return {
        "source": "tests/testthat/test.model.R",
        "theoretical_depth_ratio": {  # :4-11, expect_equal (tolerance 1.5e-8)
            "lines": "4-11",
            "cases": [
                {"CNt": 3, "cellularity": 0.5, "ploidy": 2, "expect": 1.25},
                {"CNt": 3, "cellularity": 0.8, "ploidy": 2, "expect": 1.4},
                {"CNt": 4, "cellularity": 0.8, "ploidy": 4, "expect": 1},
            ],
        },
        "theoretical_baf": {  # :13-20
            "lines": "13-20",
            "cases": [
                {"CNt": 4, "B": 2, "cellularity": 0.5, "expect": 0.5, "round": None},
                {"CNt": 3, "B": 1, "cellularity": 0.8, "expect": 0.357, "round": 3},
                {"CNt": 4, "B": 0, "cellularity": 0.8, "expect": 0.056, "round": 3},
            ],
        },
        "theoretical_mufreq": {  # :22-29
            "lines": "22-29",
            "cases": [
                {"CNt": 4, "Mt": 2, "cellularity": 0.5, "expect": 0.333, "round": 3},
                {"CNt": 3, "Mt": 1, "cellularity": 0.8, "expect": 0.286, "round": 3},
                {"CNt": 4, "Mt": 0, "cellularity": 0.8, "expect": 0, "round": None},
            ],
        },
        "baf_types_matrix": {  # :31-45  baf.types.matrix(0, 3, 2)
            "lines": "31-45",
            "args": {"CNt_min": 0, "CNt_max": 3, "CNn": 2},
            "CNn": [2, 2, 2, 2, 2, 2],
            "CNt": [0, 1, 2, 2, 3, 3],
            "B": [0, 0, 0, 1, 0, 1],
        },
        "mufreq_types_matrix": {  # :47-55  mufreq.types.matrix(0, 3, 2)
            "lines": "47-55",
            "args": {"CNt_min": 0, "CNt_max": 3, "CNn": 2},
            "CNn": [2] * 10,
            "CNt": [0, 1, 1, 2, 2, 2, 3, 3, 3, 3],
            "Mt": [0, 0, 1, 0, 1, 2, 0, 1, 2, 3],
        },
    }


def main():
    for name, fn in (("cp_example.json", cp_example), ("vignette_rows.json", vignette_rows),
                     ("testthat_model.json", testthat_model)):
        with open(os.path.join(HERE, name), "w") as f:
            json.dump(fn(), f, indent=1)
        print("wrote", name)


if __name__ == "__main__":
    main()
