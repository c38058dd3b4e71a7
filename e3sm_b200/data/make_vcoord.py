"""Convert HOMME's hybrid-coordinate ASCII tables into vcoord.npz.

Run once in the build container (reads /root/reference, which does not exist on the GPU box):
    python e3sm_b200/data/make_vcoord.py
File format: hybvcoord_mod.F90:83-117 (count, hyai values, count, hybi values; same for mid levels).
"""
import sys, os
import numpy as np

SRC = "/root/reference/components/homme/test/vcoord"
TABLES = {26: ("cami-26.ascii", "camm-26.ascii"),
          72: ("acme-72i.ascii", "acme-72m.ascii"),
          128: ("sabi-128.ascii", "sabm-128.ascii")}


def read_two(path):
    # list-directed Fortran read: text after '!' on a line is a comment
    toks = ' '.join(l.split('!')[0] for l in open(path)).split()
    n = int(toks[0]); a = np.array([float(t) for t in toks[1:1 + n]])
    m = int(toks[1 + n]); assert m == n
    b = np.array([float(t) for t in toks[2 + n:2 + 2 * n]])
    return a, b


def main():
    out = {}
    for nlev, (fi, fm) in TABLES.items():
        hyai, hybi = read_two(os.path.join(SRC, fi))
        hyam, hybm = read_two(os.path.join(SRC, fm))
        assert hyai.size == nlev + 1 and hyam.size == nlev
        out[f"hyai{nlev}"] = hyai; out[f"hybi{nlev}"] = hybi
        out[f"hyam{nlev}"] = hyam; out[f"hybm{nlev}"] = hybm
    np.savez(os.path.join(os.path.dirname(os.path.abspath(__file__)), "vcoord.npz"), **out)


if __name__ == "__main__":
    main()
