"""Init-time cubed-sphere geometry for HOMME's theta-l dycore (np=4): inputs of the hot path.

This is the host-side restatement of the Fortran init code that feeds ``init_elements_2d_c``
(H/src/theta-l_kokkos/theta_f2c_mod.F90:68-86).  H/ = components/homme/ in the reference.

  GLL points / Dvv        H/src/share/quadrature_mod.F90:279, derivative_mod.F90:174-209
  element corners         H/src/share/cube_mod.F90:810-846 (set_corner_coordinates), :950-962
  cube face -> sphere     H/src/share/coordinate_systems_mod.F90:308-401, 506-538
  element-local map       H/src/share/cube_mod.F90:704-769 (dmap_elementlocal), :2043-2104
  metric terms            H/src/share/cube_mod.F90:220-495 (metric_atomic)
  area correction         H/src/share/geometry_mod.F90:79-145 (set_area_correction_map2)
  mass matrices           H/src/share/mass_matrix_mod.F90:28-121
  Coriolis                H/src/share/cube_mod.F90:782-807
  DSS contributor order   H/src/share/edge_mod.F90:862-944

Arrays are returned in Fortran element layout flattened to C order, e.g. D[ie, b, a, j, i] for
Fortran ``D(i,j,a,b)``, so ``arr[ie].ravel()`` is exactly the memory the Fortran driver passes.
"""
import numpy as np

NP = 4
REARTH = 6.376e6          # physical_constants.F90:58
OMEGA = 7.292e-5          # physical_constants.F90:62
G = 9.80616
RGAS = 287.04
CP = 1005.0
P0 = 100000.0
KAPPA = RGAS / CP
TREF = 288.0
DD_PI = 3.141592653589793238462643383279


def gll():
    """np=4 Gauss-Lobatto-Legendre points and weights."""
    a = np.sqrt(1.0 / 5.0)
    pts = np.array([-1.0, -a, a, 1.0])
    wts = np.array([1.0 / 6.0, 5.0 / 6.0, 5.0 / 6.0, 1.0 / 6.0])
    return pts, wts


def dvv_matrix():
    """Dvv(i,l) as a [l][i] C array == Fortran column-major Dvv (derivative_mod.F90:190-207)."""
    x, _ = gll()
    leg = (5 * x ** 3 - 3 * x) / 2          # P_3 at the GLL points
    d = np.zeros((NP, NP))                   # d[j, i] == Fortran dvv(j,i) (row j, column i)
    for j in range(NP):
        for i in range(NP):
            if i != j:
                d[j, i] = (1.0 / (x[i] - x[j])) * leg[i] / leg[j]
    d[NP - 1, NP - 1] = NP * (NP - 1) / 4.0
    d[0, 0] = -NP * (NP - 1) / 4.0
    # Fortran memory of dvv(j,i): offset j + 4*i  -> C array [i][j]
    return np.ascontiguousarray(d.T)


def _face_to_sphere(x, y, face):
    """unit_face_based_cube_to_unit_sphere (coordinate_systems_mod.F90:308-401); x,y on the unit cube."""
    r = np.sqrt(1.0 + x * x + y * y)
    one = np.ones_like(x)
    lat = np.empty_like(x)
    lon = np.empty_like(x)
    for f in range(1, 7):
        m = face == f
        if not m.any():
            continue
        xx, yy, rr, o = x[m], y[m], r[m], one[m]
        if f == 1:
            la, lo = np.arcsin(yy / rr), np.arctan2(xx, o)
        elif f == 2:
            la, lo = np.arcsin(yy / rr), np.arctan2(o, -xx)
        elif f == 3:
            la, lo = np.arcsin(yy / rr), np.arctan2(-xx, -o)
        elif f == 4:
            la, lo = np.arcsin(yy / rr), np.arctan2(-o, xx)
        elif f == 5:
            big = (np.abs(yy) > 1e-9) | (np.abs(xx) > 1e-9)
            lo = np.where(big, np.arctan2(xx, yy), 0.0)
            la = np.arcsin(-o / rr)
        else:
            big = (np.abs(yy) > 1e-9) | (np.abs(xx) > 1e-9)
            lo = np.where(big, np.arctan2(xx, -yy), 0.0)
            la = np.arcsin(o / rr)
        lat[m], lon[m] = la, lo
    lon = np.where(lon < 0.0, lon + 2 * DD_PI, lon)
    return lat, lon


def _sphere_tri_area(v1, v2, v3):
    """sphere_tri_area (coordinate_systems_mod.F90:770-830), vectorised over leading axis."""
    al = np.arccos(np.sum(v2 * v3, -1))
    bl = np.arccos(np.sum(v3 * v1, -1))
    cl = np.arccos(np.sum(v1 * v2, -1))
    sina = np.sin((bl + cl - al) / 2)
    sinb = np.sin((al + cl - bl) / 2)
    sinc = np.sin((al + bl - cl) / 2)
    sins = np.sin((al + bl + cl) / 2)
    a = np.sqrt((sinb * sinc) / (sins * sina))
    b = np.sqrt((sina * sinc) / (sins * sinb))
    c = np.sqrt((sina * sinb) / (sins * sinc))
    a1, b1, c1 = 2 * np.arctan(a), 2 * np.arctan(b), 2 * np.arctan(c)
    ca = (a > b) & (a > c)
    cb = ~ca & (b > c)
    cc = ~ca & ~cb
    a1 = np.where(ca, -2 * np.arctan(1 / a), a1)
    b1 = np.where(cb, -2 * np.arctan(1 / b), b1)
    c1 = np.where(cc, -2 * np.arctan(1 / c), c1)
    return a1 + b1 + c1


def _integer_labels(ne, face, ie, je):
    """Exact integer cube-surface label of every GLL point: shared nodes get identical labels.

    Face orientation follows cart2spherical (coordinate_systems_mod.F90:403-496):
    face1 (X,Y,Z)~(1,x,y); 2 ~(-x,1,y); 3 ~(-1,-x,y); 4 ~(x,-1,y); 5 ~(y,x,-1); 6 ~(-y,x,1).
    """
    N = 3 * ne
    i = np.arange(NP)
    a = 3 * ie[:, None, None] + i[None, None, :]        # [e, j, i] -> x index
    b = 3 * je[:, None, None] + i[None, :, None]        # y index
    a = np.broadcast_to(a, (ie.size, NP, NP))
    b = np.broadcast_to(b, (ie.size, NP, NP))
    sx, sy = 2 * a - N, 2 * b - N
    f = face[:, None, None]
    Nn = np.full_like(sx, N)
    X = np.select([f == 1, f == 2, f == 3, f == 4, f == 5, f == 6], [Nn, -sx, -Nn, sx, sy, -sy])
    Y = np.select([f == 1, f == 2, f == 3, f == 4, f == 5, f == 6], [sx, Nn, -sx, -Nn, sx, sx])
    Z = np.select([f == 1, f == 2, f == 3, f == 4, f == 5, f == 6], [sy, sy, sy, sy, -Nn, Nn])
    M = 2 * N + 1
    lab = ((X + N).astype(np.int64) * M + (Y + N)) * M + (Z + N)
    return lab.reshape(ie.size, NP * NP)


def build_dss_tables(labels):
    """From per-point node labels [nelem,16] build

    node_id[nelem,16]  dense unique-node ids,
    dss_src[nelem,16,3] neighbour sources (elem*16+pt, -1 pad) in the Fortran unpack order
                        (edge_mod.F90:888-941: edges S,N,W,E inside one i-loop, then corners).
    """
    nelem = labels.shape[0]
    flat = labels.ravel()
    uniq, node_id = np.unique(flat, return_inverse=True)
    node_id = node_id.reshape(nelem, 16)
    order = np.argsort(node_id.ravel(), kind="stable")
    sorted_ids = node_id.ravel()[order]
    start = np.searchsorted(sorted_ids, np.arange(uniq.size))
    count = np.diff(np.append(start, sorted_ids.size))
    assert count.max() <= 4
    grp = -np.ones((uniq.size, 4), dtype=np.int64)
    for s in range(4):
        m = count > s
        grp[m, s] = order[start[m] + s]
    # all contributors of each (e,p), self removed
    cand = grp[node_id]                                   # [e,16,4]
    self_idx = (np.arange(nelem)[:, None] * 16 + np.arange(16)[None, :])[:, :, None]
    cand = np.where(cand == self_idx, -1, cand)
    # edge neighbours from an interior edge point: S p=1, N p=13, W p=4, E p=7
    def other(p):
        c = cand[:, p, :]
        return (c.max(axis=1)) // 16
    nS, nN, nW, nE = other(1), other(13), other(4), other(7)
    celem = np.where(cand >= 0, cand // 16, -1)           # [e,16,4]
    prio = np.full(cand.shape, 9, dtype=np.int64)
    prio[cand >= 0] = 2
    pts = np.arange(16)
    pi, pj = pts % 4, pts // 4
    onS, onN, onW, onE = pj == 0, pj == 3, pi == 0, pi == 3
    isS = (celem == nS[:, None, None]) & onS[None, :, None] & (cand >= 0)
    isN = (celem == nN[:, None, None]) & onN[None, :, None] & (cand >= 0)
    isW = (celem == nW[:, None, None]) & onW[None, :, None] & (cand >= 0)
    isE = (celem == nE[:, None, None]) & onE[None, :, None] & (cand >= 0)
    # order at corners: SW: S,W ; SE: E,S ; NW: N,W ; NE: N,E   (see module docstring)
    first = np.zeros(cand.shape, bool)
    second = np.zeros(cand.shape, bool)
    first |= isN                       # N is first wherever it applies (NW, NE, N edge)
    first |= isS & ~onE[None, :, None]  # S first on S edge and SW corner
    first |= isE & onS[None, :, None]   # E first at SE corner
    first |= isE & ~onS[None, :, None] & ~onN[None, :, None]  # plain E edge
    first |= isW & ~onS[None, :, None] & ~onN[None, :, None]  # plain W edge
    second |= isW & (onS | onN)[None, :, None]
    second |= isS & onE[None, :, None]
    second |= isE & onN[None, :, None]
    prio[first] = 0
    prio[second & ~first] = 1
    idx = np.argsort(prio, axis=2, kind="stable")
    cand_sorted = np.take_along_axis(cand, idx, axis=2)
    prio_sorted = np.take_along_axis(prio, idx, axis=2)
    assert (cand_sorted[:, :, 3] == -1).all()
    dss_src = cand_sorted[:, :, :3].astype(np.int32)
    # sanity: no two valid entries share priority 0 or 1
    v = prio_sorted[:, :, :3]
    assert not ((v[:, :, 0] == v[:, :, 1]) & (v[:, :, 0] < 2)).any()
    nbrs = np.stack([nS, nN, nW, nE], axis=1).astype(np.int32)
    return node_id.astype(np.int32), dss_src, nbrs


def dss_numpy(field, dss_src):
    """Reference DSS sum for tests: field [nelem, ..., 16] -> summed copy (own + neighbours in order)."""
    nelem = field.shape[0]
    f = field.reshape(nelem, -1, 16)
    out = f.copy()
    for j in range(3):
        s = dss_src[:, :, j]
        valid = s >= 0
        se, sp = np.where(valid, s // 16, 0), np.where(valid, s % 16, 0)
        contrib = f[se, :, sp]                      # [nelem,16,L]
        contrib = np.where(valid[:, :, None], contrib, 0.0)
        out = out + np.transpose(contrib, (0, 2, 1))
    return out.reshape(field.shape)


def _dmap_elementlocal(c3d, a, b, lon, lat):
    """dmap_elementlocal (cube_mod.F90:704-769). c3d [e,4,3]; a,b scalars; lon,lat [e] -> D [e,2,2] (D[e,i,j])."""
    q = np.array([(1 - a) * (1 - b), (1 + a) * (1 - b), (1 + a) * (1 + b), (1 - a) * (1 + b)]) / 4.0
    xx = np.einsum("eci,c->ei", c3d, q)
    r = np.sqrt(xx[:, 0] ** 2 + xx[:, 1] ** 2 + xx[:, 2] ** 2)
    sl, st, cl, ct = np.sin(lon), np.sin(lat), np.cos(lon), np.cos(lat)
    z, o = np.zeros_like(sl), np.ones_like(sl)
    D1 = np.stack([np.stack([-sl, cl, z], -1), np.stack([z, z, o], -1)], 1)            # [e,2,3]
    D2 = np.stack([np.stack([sl ** 2 * ct ** 2 + st ** 2, -sl * cl * ct ** 2, -cl * st * ct], -1),
                   np.stack([-sl * cl * ct ** 2, cl ** 2 * ct ** 2 + st ** 2, -sl * st * ct], -1),
                   np.stack([-cl * st, -sl * st, ct], -1)], 1)                          # [e,3,3]
    dd = np.array([[-1 + b, -1 + a], [1 - b, -1 - a], [1 + b, 1 + a], [-1 - b, 1 - a]]) / 4.0
    D3 = np.einsum("eci,cj->eij", c3d, dd)                                              # [e,3,2]
    D4 = np.einsum("eik,ekj->eij", D2, D3)
    D = np.einsum("eik,ekj->eij", D1, D4)
    return D / r[:, None, None]


class CubedSphere:
    """All per-element inputs of the hot path for an ne x ne x 6 cubed sphere, np=4.

    ``order``: optional permutation of global element numbers (0-based, cube_mod numbering) giving
    the storage order of elements (e.g. the space-filling curve).  Attributes use the Fortran
    memory layout per element (see module docstring).
    """

    def __init__(self, ne, order=None, rearth=REARTH, hypervis_scaling=0.0):
        self.ne = ne
        self.hypervis_scaling = float(hypervis_scaling)
        nelem = 6 * ne * ne
        gnum = np.arange(nelem) if order is None else np.asarray(order, dtype=np.int64)
        assert np.array_equal(np.sort(gnum), np.arange(nelem))
        self.nelem = nelem
        self.gnum = gnum
        self.rearth = rearth
        pts, wts = gll()
        self.dvv = dvv_matrix()
        self.mp = np.outer(wts, wts)                       # mp[j,i] = w_i w_j (symmetric)
        face = gnum // (ne * ne) + 1                       # convert_gbl_index cube_mod.F90:950-962
        ie = gnum % ne
        je = gnum // ne - (face - 1) * ne
        self.face, self.ie_face, self.je_face = face, ie, je
        dx = (0.25 * DD_PI - (-0.25 * DD_PI)) / ne
        sx = -0.25 * DD_PI + ie * dx
        sy = -0.25 * DD_PI + je * dx
        cx = np.stack([sx, sx + dx, sx + dx, sx], 1)       # corners(1..4) cube_mod.F90:836-843
        cy = np.stack([sy, sy, sy + dx, sy + dx], 1)
        f4 = np.repeat(face[:, None], 4, 1)
        lat, lon = _face_to_sphere(np.tan(cx), np.tan(cy), f4)
        c3d = np.stack([np.cos(lat) * np.cos(lon), np.cos(lat) * np.sin(lon), np.sin(lat)], -1)  # [e,4,3]
        self.corners3D = c3d
        # spherep via the element-local map (cube_mod.F90:2070-2104)
        A, B = np.meshgrid(pts, pts, indexing="xy")        # A[j,i]=pts[i], B[j,i]=pts[j]
        q = np.stack([(1 - A) * (1 - B), (1 + A) * (1 - B), (1 + A) * (1 + B), (1 - A) * (1 + B)], 0) / 4.0
        xx = np.einsum("eci,cjk->ejki", c3d, q)            # [e,j,i,3]
        r = np.sqrt(xx[..., 0] ** 2 + xx[..., 1] ** 2 + xx[..., 2] ** 2)
        cart = xx / r[..., None]
        self.cart = cart.reshape(nelem, 16, 3)
        slat = np.arcsin(cart[..., 2])
        slon = np.arctan2(cart[..., 1], cart[..., 0])
        slon = np.where(slon < 0, slon + 2 * DD_PI, slon)
        slon = np.where(np.abs(np.abs(slat) - DD_PI / 2) >= 1e-9, slon, 0.0)
        self.lat = slat.reshape(nelem, 16)
        self.lon = slon.reshape(nelem, 16)
        # D at every GLL point
        D = np.empty((nelem, NP, NP, 2, 2))                # [e,j,i,row,col] = D(i,j,row,col)
        for j in range(NP):
            for i in range(NP):
                D[:, j, i] = _dmap_elementlocal(c3d, pts[i], pts[j], slon[:, j, i], slat[:, j, i])
        # area correction alpha (geometry_mod.F90:104-131)
        detD = D[..., 0, 0] * D[..., 1, 1] - D[..., 0, 1] * D[..., 1, 0]
        metdet0 = np.abs(detD)
        area_elem = _sphere_tri_area(c3d[:, 0], c3d[:, 1], c3d[:, 2]) + _sphere_tri_area(c3d[:, 0], c3d[:, 2], c3d[:, 3])
        area_num = np.zeros(nelem)
        sum_w = np.zeros(nelem)
        for j in range(NP):
            for i in range(NP):
                area_num = area_num + wts[i] * wts[j] * metdet0[:, j, i]
        for j in range(1, NP - 1):
            for i in range(1, NP - 1):
                sum_w = sum_w + wts[i] * wts[j] * metdet0[:, j, i]
        alpha = 1.0 + (area_elem - area_num) / sum_w
        self.alpha = alpha
        # metric_atomic (cube_mod.F90:250-286, 479-494)
        met11 = D[..., 0, 0] * D[..., 0, 0] + D[..., 1, 0] * D[..., 1, 0]
        met12 = D[..., 0, 0] * D[..., 0, 1] + D[..., 1, 0] * D[..., 1, 1]
        met22 = D[..., 0, 1] * D[..., 0, 1] + D[..., 1, 1] * D[..., 1, 1]
        Dinv = np.empty_like(D)
        Dinv[..., 0, 0] = D[..., 1, 1] / detD
        Dinv[..., 0, 1] = -D[..., 0, 1] / detD
        Dinv[..., 1, 0] = -D[..., 1, 0] / detD
        Dinv[..., 1, 1] = D[..., 0, 0] / detD
        metdet = np.abs(detD)
        rmetdet = 1.0 / np.abs(detD)
        metinv = np.empty_like(D)
        metinv[..., 0, 0] = met22 / (detD * detD)
        metinv[..., 0, 1] = -met12 / (detD * detD)
        metinv[..., 1, 0] = -met12 / (detD * detD)
        metinv[..., 1, 1] = met11 / (detD * detD)
        # tensor for the resolution-aware hyperviscosity, from the UNCORRECTED D and metinv (cube_mod.F90:313-428
        # runs before the area correction :458-494):  V = (D E) diag(lamStar_i^2 eig_i) (D E)^T with
        # metinv = E diag(eig) E^T and lamStar_i = eig_i^(-hypervis_scaling/4) rearth^2
        Vt = None
        if self.hypervis_scaling != 0.0:
            eig, E = np.linalg.eigh(metinv)                              # [e,j,i,2], [e,j,i,2,2]
            lam = eig ** (1.0 - self.hypervis_scaling / 2.0) * rearth ** 4
            DE = np.einsum("ejiab,ejibc->ejiac", D, E)
            Vt = np.einsum("ejiac,ejic,ejibc->ejiab", DE, lam, DE)      # [e,j,i,a,b] = tensorVisc(i,j,a,b)
        sa = np.sqrt(alpha)[:, None, None, None, None]
        al = alpha[:, None, None]
        inner = (slice(None), slice(1, NP - 1), slice(1, NP - 1))
        D[inner] = D[inner] * sa
        Dinv[inner] = Dinv[inner] / sa
        metdet[inner] = metdet[inner] * al
        rmetdet[inner] = 1.0 / metdet[inner]
        metinv[inner] = metinv[inner] / al[..., None, None]
        # to Fortran memory: X(i,j,a,b) -> [e,b,a,j,i]
        tr = lambda X: np.ascontiguousarray(np.transpose(X, (0, 4, 3, 1, 2))).reshape(nelem, 2, 2, 16)
        self.D, self.Dinv, self.metinv = tr(D), tr(Dinv), tr(metinv)
        self.metdet = metdet.reshape(nelem, 16)
        self.rmetdet = rmetdet.reshape(nelem, 16)
        self.fcor = 2.0 * OMEGA * np.sin(self.lat)                        # cube_mod.F90:803
        self.spheremp = self.mp.reshape(1, 16) * self.metdet             # mass_matrix_mod.F90:98
        # vec_sphere2cart(np,np,3,2) -> [e,comp2,comp3,p]  (cube_mod.F90:188-198)
        la, lo = self.lat, self.lon
        z = np.zeros_like(la)
        v2c = np.stack([np.stack([-np.sin(lo), np.cos(lo), z], 1),
                        np.stack([-np.sin(la) * np.cos(lo), -np.sin(la) * np.sin(lo), np.cos(la)], 1)], 1)
        self.vec_sph2cart = np.ascontiguousarray(v2c)                    # [e,2,3,16]
        self.tensorvisc = np.zeros((nelem, 2, 2, 16))                    # hypervis_scaling=0 configs
        # DSS tables + rspheremp
        labels = _integer_labels(ne, face, ie, je)
        self.node_id, self.dss_src, self.edge_nbrs = build_dss_tables(labels)
        self.nnodes = int(self.node_id.max()) + 1
        self.rspheremp = 1.0 / dss_numpy(self.spheremp, self.dss_src)    # mass_matrix_mod.F90:100-114
        if Vt is not None:
            # dss_hvtensor (global_norms_mod.F90:415-486): mass-weighted DSS of every component, then the
            # bilinear function of the four (now continuous) corner values inside each element
            V = Vt.reshape(nelem, 16, 2, 2)
            x = np.tile(pts, NP)                                         # x(p) = pts[i], y(p) = pts[j], p = j*4 + i
            y = np.repeat(pts, NP)
            out = np.zeros((nelem, 2, 2, 16))                            # [e,b,a,p]
            for a in range(2):
                for b in range(2):
                    z = dss_numpy(V[:, :, a, b] * self.spheremp, self.dss_src) * self.rspheremp
                    sw, se, nw, ne_ = z[:, 0:1], z[:, 3:4], z[:, 12:13], z[:, 15:16]
                    out[:, b, a, :] = 0.25 * ((1.0 - x) * (1.0 - y) * sw + (1.0 - x) * (y + 1.0) * nw
                                              + (x + 1.0) * (1.0 - y) * se + (x + 1.0) * (y + 1.0) * ne_)
            self.tensorvisc = np.ascontiguousarray(out)

    def local_view(self, e0, e1):
        """Per-element arrays of the storage slice [e0:e1) (what one rank of the SFC partition owns)."""
        class _View:
            pass
        v = _View()
        v.nelem, v.rearth, v.ne = e1 - e0, self.rearth, self.ne
        for name in ("lat", "lon", "D", "Dinv", "metinv", "metdet", "rmetdet", "fcor", "spheremp",
                     "rspheremp", "gnum", "cart"):
            setattr(v, name, getattr(self, name)[e0:e1])
        return v

    def sphere_area(self):
        return float(self.spheremp.sum())
