"""Evaluate a spectral-element field at arbitrary points of the sphere (host side, for diagnostics and the
resolution-convergence test): element search by inverting the element-local map (cube_mod.F90:2070-2104, the
bilinear combination of the corner vectors projected to the sphere) and np=4 Lagrange interpolation."""
import numpy as np
from scipy.spatial import cKDTree

from .geometry import gll


def _lagrange(pts, x):
    """L[i] of the np Lagrange polynomials on pts at the points x: [n, np]"""
    n = len(pts)
    L = np.ones((x.shape[0], n))
    for i in range(n):
        for m in range(n):
            if m != i:
                L[:, i] *= (x - pts[m]) / (pts[i] - pts[m])
    return L


def locate(grid, xyz, ncand=6):
    """element index and reference coordinates (a,b) in [-1,1]^2 of every unit vector xyz[n,3]"""
    c = grid.corners3D                                          # [e,4,3]
    centre = c.mean(axis=1)
    centre /= np.linalg.norm(centre, axis=1, keepdims=True)
    _, cand = cKDTree(centre).query(xyz, k=ncand)
    n = xyz.shape[0]
    elem = np.full(n, -1)
    ab = np.zeros((n, 2))
    best = np.full(n, np.inf)
    for j in range(ncand):
        e = cand[:, j]
        cc = c[e]                                               # [n,4,3]
        a = np.zeros(n); b = np.zeros(n)
        for _ in range(12):                                     # Gauss-Newton on normalise(sum q_c corner_c) = xyz
            q = np.stack([(1 - a) * (1 - b), (1 + a) * (1 - b), (1 + a) * (1 + b), (1 - a) * (1 + b)], 1) / 4
            qa = np.stack([-(1 - b), (1 - b), (1 + b), -(1 + b)], 1) / 4
            qb = np.stack([-(1 - a), -(1 + a), (1 + a), (1 - a)], 1) / 4
            x = np.einsum("nc,nci->ni", q, cc)
            r = np.linalg.norm(x, axis=1, keepdims=True)
            u = x / r
            xa = np.einsum("nc,nci->ni", qa, cc); xb = np.einsum("nc,nci->ni", qb, cc)
            ua = (xa - u * np.sum(u * xa, 1, keepdims=True)) / r
            ub = (xb - u * np.sum(u * xb, 1, keepdims=True)) / r
            res = xyz - u
            g11, g12, g22 = np.sum(ua * ua, 1), np.sum(ua * ub, 1), np.sum(ub * ub, 1)
            r1, r2 = np.sum(ua * res, 1), np.sum(ub * res, 1)
            det = g11 * g22 - g12 * g12
            a = a + (g22 * r1 - g12 * r2) / det
            b = b + (g11 * r2 - g12 * r1) / det
        out = np.maximum(np.abs(a), np.abs(b))
        take = out < best
        elem[take], best[take] = e[take], out[take]
        ab[take, 0], ab[take, 1] = a[take], b[take]
    assert (best <= 1 + 1e-8).all(), "point outside every candidate element"
    return elem, np.clip(ab, -1.0, 1.0)


def evaluate(grid, field, xyz):
    """field[e,16] (C0, np=4) at the unit vectors xyz[n,3]"""
    pts, _ = gll()
    e, ab = locate(grid, xyz)
    La, Lb = _lagrange(pts, ab[:, 0]), _lagrange(pts, ab[:, 1])      # [n,4]
    f = field[e].reshape(-1, 4, 4)                                    # [n,j,i]
    return np.einsum("nj,nji,ni->n", Lb, f, La)
