"""An independent derivation of the per-element geometry, in 40-digit arithmetic, against
e3sm_b200/geometry.py.  Nothing here shares code with geometry.py except the element corner coordinates:

  * D (cube_mod.F90:704-769, Dmap): centred differences of the element-local map (a,b) -> (lon,lat),
    D(1,:) = cos(lat) dlon/d(a,b), D(2,:) = dlat/d(a,b), instead of the analytic chain of three matrices;
  * metdet, Dinv, metinv from that D (cube_mod.F90:262-316);
  * the epsilon-bubble area correction (cube_mod.F90:458-494, geometry_mod.F90:79-145) from the spherical excess
    of the element (Girard) instead of the two-triangle l'Huilier formula;
  * spheremp, rspheremp (mass_matrix_mod.F90:28-121) with the DSS done by matching node COORDINATES, not the
    integer labels / gather table geometry.py builds;
  * Dvv (derivative_mod.F90:190-207) from differentiating the Lagrange polynomials symbolically.
"""
import numpy as np
import pytest
from mpmath import mp, mpf, sqrt, atan2, asin, cos, acos, matrix

mp.dps = 40


def _unit(v):
    r = sqrt(v[0] ** 2 + v[1] ** 2 + v[2] ** 2)
    return [v[0] / r, v[1] / r, v[2] / r]


def _point(corners, a, b):
    q = [(1 - a) * (1 - b) / 4, (1 + a) * (1 - b) / 4, (1 + a) * (1 + b) / 4, (1 - a) * (1 + b) / 4]
    x = [sum(q[c] * corners[c][i] for c in range(4)) for i in range(3)]
    return _unit(x)


def _lonlat(x):
    return atan2(x[1], x[0]), asin(x[2])


def _D_fd(corners, a, b):
    h = mpf(10) ** (-12)
    lon0, lat0 = _lonlat(_point(corners, a, b))
    D = matrix(2, 2)
    for j, (da, db) in enumerate(((h, 0), (0, h))):
        lp, tp = _lonlat(_point(corners, a + da, b + db))
        lm, tm = _lonlat(_point(corners, a - da, b - db))
        dl = lp - lm
        if dl > mp.pi:
            dl -= 2 * mp.pi
        if dl < -mp.pi:
            dl += 2 * mp.pi
        D[0, j] = cos(lat0) * dl / (2 * h)
        D[1, j] = (tp - tm) / (2 * h)
    return D


def _angle_at(p, q, r):
    """interior angle at p of the spherical polygon corner q-p-r"""
    def tangent(a, b):
        d = sum(a[i] * b[i] for i in range(3))
        return _unit([b[i] - d * a[i] for i in range(3)])
    t1, t2 = tangent(p, q), tangent(p, r)
    return acos(sum(t1[i] * t2[i] for i in range(3)))


@pytest.fixture(scope="module")
def grid():
    from e3sm_b200.geometry import CubedSphere
    return CubedSphere(3)


def test_dvv_from_lagrange_polynomials(grid):
    pts = [-mpf(1), -1 / sqrt(mpf(5)), 1 / sqrt(mpf(5)), mpf(1)]
    for i in range(4):          # d/dx of the i-th Lagrange polynomial at pts[l]
        for l in range(4):
            s = mpf(0)
            for m in range(4):
                if m == i:
                    continue
                t = 1 / (pts[i] - pts[m])
                for n in range(4):
                    if n != i and n != m:
                        t *= (pts[l] - pts[n]) / (pts[i] - pts[n])
                s += t
            # geometry.dvv[l, i] = Dvv(i,l) = derivative of basis function i at node l
            assert abs(float(s) - grid.dvv[l, i]) < 2e-15, (i, l)
    w = [mpf(1) / 6, mpf(5) / 6, mpf(5) / 6, mpf(1) / 6]
    for j in range(4):
        for i in range(4):
            assert abs(float(w[i] * w[j]) - grid.mp[j, i]) < 3e-16


@pytest.mark.parametrize("e", [0, 4, 13, 22, 31, 40, 53])
def test_metric_terms_against_finite_differences(grid, e):
    pts = [-mpf(1), -1 / sqrt(mpf(5)), 1 / sqrt(mpf(5)), mpf(1)]
    w = [mpf(1) / 6, mpf(5) / 6, mpf(5) / 6, mpf(1) / 6]
    corners = [[mpf(float(x)) for x in grid.corners3D[e, c]] for c in range(4)]
    D = [[_D_fd(corners, pts[i], pts[j]) for i in range(4)] for j in range(4)]       # [j][i]
    det = [[D[j][i][0, 0] * D[j][i][1, 1] - D[j][i][0, 1] * D[j][i][1, 0] for i in range(4)] for j in range(4)]
    # area correction: geometric area by the spherical excess, numerical area by GLL quadrature
    ang = sum(_angle_at(corners[c], corners[(c + 1) % 4], corners[(c - 1) % 4]) for c in range(4))
    area = ang - 2 * mp.pi
    num = sum(w[i] * w[j] * abs(det[j][i]) for i in range(4) for j in range(4))
    inner = sum(w[i] * w[j] * abs(det[j][i]) for i in (1, 2) for j in (1, 2))
    alpha = 1 + (area - num) / inner
    assert abs(float(alpha) - grid.alpha[e]) < 1e-12
    for j in range(4):
        for i in range(4):
            p = j * 4 + i
            a = alpha if (i in (1, 2) and j in (1, 2)) else mpf(1)
            Dc = D[j][i] * sqrt(a)
            dt = Dc[0, 0] * Dc[1, 1] - Dc[0, 1] * Dc[1, 0]
            Dinv = matrix([[Dc[1, 1], -Dc[0, 1]], [-Dc[1, 0], Dc[0, 0]]]) / dt
            met = Dc.T * Dc
            metinv = met ** -1
            scale = float(abs(Dc[0, 0]) + abs(Dc[1, 1]))
            for r in range(2):
                for c in range(2):                      # geometry layout [e, col, row, p] = X(i,j,row,col)
                    assert abs(float(Dc[r, c]) - grid.D[e, c, r, p]) < 1e-11 * scale, ("D", p, r, c)
                    assert abs(float(Dinv[r, c]) - grid.Dinv[e, c, r, p]) < 1e-11 / scale, ("Dinv", p, r, c)
                    assert abs(float(metinv[r, c]) - grid.metinv[e, c, r, p]) < 1e-10 / scale ** 2, ("metinv", p)
            assert abs(float(abs(dt)) - grid.metdet[e, p]) < 1e-11 * float(abs(dt))
            assert abs(float(w[i] * w[j] * abs(dt)) - grid.spheremp[e, p]) < 1e-11 * grid.spheremp[e, p]
            lon, lat = _lonlat(_point(corners, pts[i], pts[j]))
            assert abs(float(lat) - grid.lat[e, p]) < 1e-13
            assert abs(float(2 * mpf("7.292e-5") * mp.sin(lat)) - grid.fcor[e, p]) < 1e-18


def test_rspheremp_by_coordinate_matching(grid):
    """1 / (sum of spheremp over the elements sharing a node), nodes identified by their coordinates"""
    key = np.round(grid.cart.reshape(-1, 3) * 1e9).astype(np.int64)
    _, inv, cnt = np.unique(key, axis=0, return_inverse=True, return_counts=True)
    assert sorted(set(cnt.tolist())) == [1, 2, 3, 4]                     # interior, edge, cube vertex, element corner
    assert inv.max() + 1 == grid.nnodes == 6 * 9 * 9 + 2                 # Euler: 6 ne^2 (np-1)^2 + 2
    tot = np.zeros(inv.max() + 1)
    np.add.at(tot, inv.ravel(), grid.spheremp.ravel())
    r = (1.0 / tot)[inv.ravel()].reshape(grid.spheremp.shape)
    assert np.abs(r / grid.rspheremp - 1).max() < 1e-14
    assert abs(grid.spheremp.sum() / (4 * np.pi) - 1) < 1e-13            # the corrected areas add up to the sphere
