"""HOMME's space-filling-curve element decomposition (one rank <-> one GPU).

Restates, for the cubed sphere:
  Factor / IsFactorable        H/src/share/spacecurve_mod.F90:874-969   (ne = 2^n 3^m 5^p)
  GenSpaceCurve / map          H/src/share/spacecurve_mod.F90:971-1020
  hilbert / PeanoM / Cinco     H/src/share/spacecurve_mod.F90:717-803, 540-715, 73-537
                               (sub-cell templates in data/sfc_tables.json, tools/make_sfc_tables.py)
  IncrementCurve               H/src/share/spacecurve_mod.F90:805-817
  non-factorable ne            H/src/share/cube_mod.F90:1036-1102 (nearest element of a 2^k mesh)
  face chaining 1,2,6,4,5,3    H/src/share/cube_mod.F90:1108-1164
  genspacepart                 H/src/share/spacecurve_mod.F90:1198-1252
H/ = components/homme/ in the reference.  Global element numbers are cube_mod's
(face-1)*ne*ne + (j-1)*ne + i, 0-based here.
"""
import json
import os

import numpy as np

_TABLES = None


def _tables():
    global _TABLES
    if _TABLES is None:
        path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "sfc_tables.json")
        _TABLES = {int(k): v for k, v in json.load(open(path)).items()}
    return _TABLES


def factor(num):
    """Factors of num in HOMME's order (all 2s, then 3s, then 5s); None if not of that form."""
    f, tmp = [], num
    for q in (2, 3, 5):
        while tmp % q == 0:
            f.append(q)
            tmp //= q
    return f if tmp == 1 else None


def is_factorable(n):
    return factor(n) is not None


def gen_space_curve(n):
    """Mesh[i, j] (0-based) = visit order of cell (i, j) on the n x n face curve."""
    fact = factor(n)
    assert fact is not None, "gen_space_curve needs n = 2^a 3^b 5^c"
    tab = _tables()
    ordered = np.zeros((n, n), dtype=np.int64)
    pos = [0, 0]
    vcnt = [0]
    if n == 1:
        return ordered

    def ev(code, ma, md, ja, jd, lma, lmd):
        if code == "ma": return ma
        if code == "ma+1": return (ma + 1) % 2
        if code == "md": return md
        if code == "-md": return -md
        if code == "lma": return lma
        if code == "lma+1": return (lma + 1) % 2
        if code == "lmd": return lmd
        if code == "-lmd": return -lmd
        if code == "ja": return ja
        if code == "jd": return jd
        raise ValueError(code)

    def gen(l, typ, ma, md, ja, jd):
        ltype = fact[l - 2] if l > 1 else None
        for c in tab[typ]:
            lma = ev(c[0], ma, md, ja, jd, None, None)
            lmd = ev(c[1], ma, md, ja, jd, lma, None)
            lja = ev(c[2], ma, md, ja, jd, lma, lmd)
            ljd = ev(c[3], ma, md, ja, jd, lma, lmd)
            if l > 1:
                gen(l - 1, ltype, lma, lmd, lja, ljd)
            else:                                   # IncrementCurve
                ordered[pos[0], pos[1]] = vcnt[0]
                vcnt[0] += 1
                pos[lja] += ljd

    level = len(fact)
    gen(level, fact[level - 1], 0, 1, 0, 1)
    assert vcnt[0] == n * n
    return ordered


def face_mesh(ne):
    """Mesh(i,j) of cube_mod.F90:1031-1102, including the non-factorable fallback."""
    if is_factorable(ne):
        return gen_space_curve(ne)
    ne2 = 2 ** int(np.ceil(np.log(float(ne)) / np.log(2.0)))
    mesh2 = gen_space_curve(ne2)
    m2map = np.zeros((ne2, ne2, 2), dtype=np.int64)          # 1-based (i,j), 0 = unmapped
    for j in range(1, ne + 1):
        for i in range(1, ne + 1):
            i2 = int(np.floor(((i - .5) / ne) * ne2 + .5 + .5))   # nint
            j2 = int(np.floor(((j - .5) / ne) * ne2 + .5 + .5))
            i2 = min(max(i2, 1), ne2)
            j2 = min(max(j2, 1), ne2)
            m2map[i2 - 1, j2 - 1] = (i, j)
    sfcij = np.zeros((ne2 * ne2, 2), dtype=np.int64)
    for j in range(ne2):
        for i in range(ne2):
            sfcij[mesh2[i, j]] = (i, j)
    mesh = np.zeros((ne, ne), dtype=np.int64)
    idx = 0
    for k in range(ne2 * ne2):
        i2, j2 = sfcij[k]
        i, j = m2map[i2, j2]
        if i != 0:
            mesh[i - 1, j - 1] = idx
            idx += 1
    assert idx == ne * ne
    return mesh


def space_curve(ne):
    """SpaceCurve index (0-based) of every global element number (cube_mod.F90:1108-1164)."""
    mesh = face_mesh(ne)
    n2 = ne * ne
    sc = np.zeros(6 * n2, dtype=np.int64)
    I, J = np.meshgrid(np.arange(ne), np.arange(ne), indexing="ij")    # 0-based i, j
    r = ne - 1
    per_face = {                       # face -> Mesh indices used for GridElem(i,j,face)
        1: (I, r - J), 2: (I, r - J), 6: (r - I, r - J), 4: (r - J, I), 5: (I, J), 3: (I, J)}
    offset = 0
    for face in (1, 2, 6, 4, 5, 3):
        a, b = per_face[face]
        g = (face - 1) * n2 + J * ne + I
        sc[g] = offset + mesh[a, b]
        offset += n2
    return sc


def genspacepart(sc, npart):
    """processor number (0-based) per element from its SpaceCurve index, remainder rule of
    spacecurve_mod.F90:1214-1243 (the first `extra` ranks get one element more)."""
    nelem = sc.size
    nelemd = nelem // npart
    extra = nelem % npart
    s1 = extra * (nelemd + 1)
    ids = sc.astype(np.int64)
    first = ids <= s1                                          # Fortran: `if (id<=s1)` on the 0-based id
    proc = np.where(first, ids // (nelemd + 1), extra + (ids - s1) // max(nelemd, 1))
    return np.minimum(proc, npart - 1).astype(np.int64)


class Partition:
    """Storage order = the space-filling curve; rank r owns the contiguous slice [start[r], start[r+1])."""

    def __init__(self, ne, npart):
        sc = space_curve(ne)
        self.ne, self.npart = ne, npart
        self.order = np.argsort(sc, kind="stable")            # storage index -> global element number
        proc_of_g = genspacepart(sc, npart)
        self.owner = proc_of_g[self.order]                     # per storage index, non-decreasing
        assert (np.diff(self.owner) >= 0).all()
        self.start = np.searchsorted(self.owner, np.arange(npart + 1))

    def local_range(self, rank):
        return int(self.start[rank]), int(self.start[rank + 1])
