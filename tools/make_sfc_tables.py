"""Dev-time generator of e3sm_b200/data/sfc_tables.json (run in the build container only).

Reads the sub-cell visit templates of HOMME's three space-filling curves (Hilbert 2x2, meandering
Peano 3x3, Cinco 5x5) from the reference, H/src/share/spacecurve_mod.F90:73-803: for every visited
sub-cell the four assignments (lma, lmd, lja, ljd) as symbolic codes of (ma, md, ja, jd).  The
recursion that consumes the tables is restated in e3sm_b200/partition.py.  The tables are data
(the definition of the published curves); /root/reference is not available on the GPU box.
"""
import json
import os
import re
import sys

SRC = "/root/reference/components/homme/src/share/spacecurve_mod.F90"
RANGES = {"5": (73, 537), "3": (540, 715), "2": (717, 803)}


def extract(lines):
    cells, cur = [], {}
    for ln in lines:
        m = re.match(r"^\s*(lma|lmd|lja|ljd)\s*=\s*(.+?)\s*$", ln)
        if not m:
            continue
        k, v = m.group(1), m.group(2).replace(" ", "")
        v = v.replace("MOD(ma+1,maxdim)", "ma+1").replace("MOD(lma+1,maxdim)", "lma+1")
        cur[k] = v
        if k == "ljd":
            cells.append([cur["lma"], cur["lmd"], cur["lja"], cur["ljd"]])
            cur = {}
    return cells


def main():
    src = open(SRC).read().split("\n")
    out = {t: extract(src[a - 1:b]) for t, (a, b) in RANGES.items()}
    assert len(out["2"]) == 4 and len(out["3"]) == 9 and len(out["5"]) == 25, {k: len(v) for k, v in out.items()}
    dst = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "e3sm_b200", "data", "sfc_tables.json")
    json.dump(out, open(dst, "w"), indent=0)
    print("wrote", dst)


if __name__ == "__main__":
    sys.exit(main())
