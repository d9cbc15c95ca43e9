"""ctypes binding of the CPU oracle (oracle/libls1_oracle.so).  TEST INFRASTRUCTURE ONLY:
imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg, never by the product."""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


class OGrid(C.Structure):
    _fields_ = [("box_min", C.c_double * 3), ("box_max", C.c_double * 3),
                ("halo_min", C.c_double * 3), ("halo_max", C.c_double * 3),
                ("cell_len", C.c_double * 3), ("cell_len_recip", C.c_double * 3),
                ("box_cells", C.c_int * 3), ("cpd", C.c_int * 3), ("cutoff", C.c_double)]


class OParams(C.Structure):
    _fields_ = [("nc", C.c_int), ("counts", C.c_void_p), ("scal", C.c_void_p), ("lj", C.c_void_p),
                ("ch", C.c_void_p), ("di", C.c_void_p), ("qu", C.c_void_p), ("mix", C.c_void_p),
                ("nmix", C.c_int), ("eps_rf", C.c_double), ("rc", C.c_double), ("rc_lj", C.c_double)]


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "libls1_oracle.so")
        if not os.path.exists(path):
            build()
        _LIB = C.CDLL(path)
        _LIB.o_cell_index.restype = C.c_uint32
        _LIB.o_c08_pairs.restype = C.c_size_t
        _LIB.o_exchange.restype = C.c_size_t
    return _LIB


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class System:
    """Component tables + cutoffs, in the flattened layout of the ref_driver dump."""

    def __init__(self, counts, scal, lj, ch, di, qu, mix, eps_rf, rc, rc_lj):
        self.counts = np.ascontiguousarray(counts, dtype=np.int64).reshape(-1, 6)
        self.scal = np.ascontiguousarray(scal, dtype=np.float64).reshape(-1, 4)
        self.lj = np.ascontiguousarray(lj, dtype=np.float64).reshape(-1, 7)
        self.ch = np.ascontiguousarray(ch, dtype=np.float64).reshape(-1, 5)
        self.di = np.ascontiguousarray(di, dtype=np.float64).reshape(-1, 7)
        self.qu = np.ascontiguousarray(qu, dtype=np.float64).reshape(-1, 7)
        self.mix = np.ascontiguousarray(mix, dtype=np.float64).ravel()
        self.eps_rf, self.rc, self.rc_lj = float(eps_rf), float(rc), float(rc_lj)
        self.params = OParams(len(self.counts), _p(self.counts), _p(self.scal), _p(self.lj), _p(self.ch),
                              _p(self.di), _p(self.qu), _p(self.mix), len(self.mix),
                              self.eps_rf, self.rc, self.rc_lj)

    @classmethod
    def from_dump(cls, d):
        return cls(d["comp_counts"], d["comp_scalars"], d["lj_sites"], d["charge_sites"], d["dipole_sites"],
                   d["quadrupole_sites"], d["mixcoeff"], float(d["epsRF"][0]), float(d["rc"][0]), float(d["rc"][1]))

    @classmethod
    def single_lj(cls, eps=1.0, sigma=1.0, mass=1.0, rc=2.5, shift6=0.0):
        return cls([[1, 0, 0, 0, 0, 0]], [[mass, 0, 0, 0]], [[0, 0, 0, mass, eps, sigma, shift6]],
                   np.zeros((0, 5)), np.zeros((0, 7)), np.zeros((0, 7)), [], 1e10, rc, rc)


def grid(box_min, box_max, rc, mode):
    g = OGrid()
    bmin = (C.c_double * 3)(*box_min)
    bmax = (C.c_double * 3)(*box_max)
    rcode = lib().o_grid_init(C.byref(g), bmin, bmax, C.c_double(rc), int(mode))
    if rcode:
        raise ValueError("region too small")
    return g


def cell_indices(g, pts):
    pts = np.ascontiguousarray(pts, dtype=np.float64).reshape(-1, 3)
    out = np.empty(len(pts), dtype=np.uint32)
    lib().o_cell_indices(C.byref(g), C.c_size_t(len(pts)), _p(pts), _p(out))
    return out


def c08_pairs(g):
    n = lib().o_c08_pairs(C.byref(g), None, C.c_size_t(0))
    out = np.empty((n, 3), dtype=np.uint32)
    lib().o_c08_pairs(C.byref(g), _p(out), C.c_size_t(n))
    return out


def lj_tables(sysm):
    S = int(sysm.counts[:, 0].sum())
    e = np.zeros((S, S)); s = np.zeros((S, S)); sh = np.zeros((S, S))
    lib().o_lj_tables(C.byref(sysm.params), _p(e), _p(s), _p(sh), S * S)
    return e, s, sh


def exchange(g, global_len, r):
    """returns (wrapped r, halo_src, halo_r)"""
    r = np.array(r, dtype=np.float64).reshape(-1, 3).copy()
    n = len(r)
    cap = 27 * n + 64
    src = np.empty(cap, dtype=np.uint32)
    hr = np.empty((cap, 3))
    gl = (C.c_double * 3)(*global_len)
    h = lib().o_exchange(C.byref(g), gl, C.c_size_t(n), _p(r), C.c_size_t(cap), _p(src), _p(hr))
    assert h <= cap
    return r, src[:h].copy(), hr[:h].copy()


def forces(sysm, g, cid, r, q):
    """returns dict(F, M, Vi, upot, virial, q(normalised), n_dist, n_pairs)"""
    cid = np.ascontiguousarray(cid, dtype=np.uint32)
    r = np.ascontiguousarray(r, dtype=np.float64).reshape(-1, 3)
    q = np.array(q, dtype=np.float64).reshape(-1, 4).copy()
    n = len(cid)
    F = np.zeros((n, 3)); M = np.zeros((n, 3)); Vi = np.zeros((n, 3))
    up = C.c_double(); vir = C.c_double()
    stats = np.zeros(2, dtype=np.uint64)
    lib().o_forces(C.byref(sysm.params), C.byref(g), C.c_size_t(n), _p(cid), _p(r), _p(q), _p(F), _p(M), _p(Vi),
                   C.byref(up), C.byref(vir), _p(stats))
    return dict(F=F, M=M, Vi=Vi, upot=up.value, virial=vir.value, q=q, n_dist=int(stats[0]), n_pairs=int(stats[1]))


def run(sysm, g, global_len, cid, r, v, q, L, F, M, steps, dt, target_T=0.0):
    cid = np.ascontiguousarray(cid, dtype=np.uint32)
    st = [np.array(a, dtype=np.float64).copy() for a in (r, v, q, L, F, M)]
    series = np.zeros((steps, 4))
    gl = (C.c_double * 3)(*global_len)
    rc = lib().o_run(C.byref(sysm.params), C.byref(g), gl, C.c_size_t(len(cid)), _p(cid), *[_p(a) for a in st],
                     int(steps), C.c_double(dt), C.c_double(target_T), _p(series))
    assert rc == 0
    return dict(r=st[0], v=st[1], q=st[2], L=st[3], F=st[4], M=st[5], series=series)


def selective_thermostat(sysm, cid, v, F, q, L, M, limit_energy, Ti):
    """numpy restatement of the reference's "explosion heuristics" molecule loop (Domain::calculateGlobalValues,
    src/Domain.cpp:266-292): U_trans = 1/2 m v^2 > limit_energy  ->  v, F *= sqrt(limit/U_trans);
    U_rot (FullMolecule::U_rot, src/molecules/FullMolecule.cpp:392-400; w = I^-1 rotateinv(q, D)) > 3 rotDOF Ti -> D, M *= sqrt(limit/U_rot).
    rotDOF = number of non-zero principal moments (Component::updateMassInertia).  Returns (v, F, L, M) copies."""
    v, F, L, M = (np.array(a, dtype=np.float64, copy=True) for a in (v, F, L, M))
    m = sysm.scal[cid, 0]
    I = sysm.scal[cid, 1:4]
    ut = 0.5 * m * (v[:, 0] * v[:, 0] + v[:, 1] * v[:, 1] + v[:, 2] * v[:, 2])
    hot = ut > limit_energy
    c = np.sqrt(limit_energy / ut[hot])
    v[hot] *= c[:, None]
    F[hot] *= c[:, None]
    w_, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    ww, xx, yy, zz = w_ * w_, x * x, y * y, z * z
    wx, wy, wz, xy, xz, yz = w_ * x, w_ * y, w_ * z, x * y, x * z, y * z
    w = np.empty_like(L)                                            # Quaternion::rotateinv (Quaternion.cpp:64-83)
    w[:, 0] = (ww + xx - yy - zz) * L[:, 0] + 2. * (xy + wz) * L[:, 1] + 2. * (xz - wy) * L[:, 2]
    w[:, 1] = 2. * (xy - wz) * L[:, 0] + (ww - xx + yy - zz) * L[:, 1] + 2. * (yz + wx) * L[:, 2]
    w[:, 2] = 2. * (xz + wy) * L[:, 0] + 2. * (yz - wx) * L[:, 1] + (ww - xx - yy + zz) * L[:, 2]
    invI = np.where(I != 0., 1. / np.where(I != 0., I, 1.), 0.)
    w *= invI
    urot = 0.5 * (I * w * w).sum(axis=1)
    rot_dof = (I != 0.).sum(axis=1)
    lim = 3.0 * rot_dof * Ti
    hot = (rot_dof > 0) & (urot > lim)
    c = np.sqrt(lim[hot] / urot[hot])
    L[hot] *= c[:, None]
    M[hot] *= c[:, None]
    return v, F, L, M


# FlopCounter multipliers (src/particleContainer/adapter/FlopCounter.cpp:19-75): kernel, macro, sum force/torque, sum macro;
# order = FlopCounter.h:138-147 (LJ, charge, charge-dipole, dipole, charge-quadrupole, dipole-quadrupole, quadrupole)
FLOP_MULT = np.array([[15, 4, 12, 2], [10, 2, 12, 2], [34, 3, 15, 2], [101, 5, 18, 3], [52, 2, 15, 2], [121, 2, 18, 2], [131, 2, 18, 2]], dtype=np.float64)


def flop_count(sysm, g, cid, r):
    """numpy restatement of the reference's FlopCounter used as the cell processor of one traversal
    (FlopCounter.cpp:196-421, FlopCounter.h:150-283).  cid, r: all molecules incl. halo copies (as for forces()).
    processCell: non-halo cells, pairs i < j; processCellPair over the C08 cell pairs (halo-halo skipped): all i x j pairs,
    kernel calls always, macro calls unless exactly one cell is halo and index(non-halo) >= index(halo) (:275-311).
    A molecule pair adds nLJ_i nLJ_j LJ calls when the CENTRE distance^2 < rcLJ^2 and the electrostatic products when < rc^2.
    Returns dict(molecule_distances, kernel[7], macro[7], flops)."""
    cid = np.asarray(cid, dtype=np.int64)
    r = np.ascontiguousarray(r, dtype=np.float64).reshape(-1, 3)
    cell = cell_indices(g, r).astype(np.int64)
    cpd = np.array(list(g.cpd), dtype=np.int64)
    order = np.argsort(cell, kind="stable")
    start = np.searchsorted(cell[order], np.arange(int(cpd.prod()) + 1))
    cnt = sysm.counts[cid, :4].astype(np.float64)                      # nLJ, nCharge, nDipole, nQuadrupole per molecule

    def is_halo(c):
        x, y, z = c % cpd[0], (c // cpd[0]) % cpd[1], c // (cpd[0] * cpd[1])
        return x == 0 or y == 0 or z == 0 or x == cpd[0] - 1 or y == cpd[1] - 1 or z == cpd[2] - 1

    rc2, rclj2 = sysm.rc * sysm.rc, sysm.rc_lj * sysm.rc_lj
    mm, kernel, macro = 0.0, np.zeros(7), np.zeros(7)

    def add(ia, ib, same, with_macro):
        nonlocal mm
        if len(ia) == 0 or len(ib) == 0:
            return
        d = r[ia][:, None, :] - r[ib][None, :, :]
        d2 = d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1] + d[..., 2] * d[..., 2]
        use = np.triu(np.ones(d2.shape, dtype=bool), 1) if same else np.ones(d2.shape, dtype=bool)
        mm += float(use.sum())
        lj = use & (d2 < rclj2)
        el = use & (d2 < rc2)
        a, b = cnt[ia], cnt[ib]
        k = np.zeros(7)
        k[0] = (np.outer(a[:, 0], b[:, 0]) * lj).sum()
        k[1] = (np.outer(a[:, 1], b[:, 1]) * el).sum()
        k[2] = ((np.outer(a[:, 1], b[:, 2]) + np.outer(a[:, 2], b[:, 1])) * el).sum()
        k[3] = (np.outer(a[:, 2], b[:, 2]) * el).sum()
        k[4] = ((np.outer(a[:, 1], b[:, 3]) + np.outer(a[:, 3], b[:, 1])) * el).sum()
        k[5] = ((np.outer(a[:, 2], b[:, 3]) + np.outer(a[:, 3], b[:, 2])) * el).sum()
        k[6] = (np.outer(a[:, 3], b[:, 3]) * el).sum()
        kernel[:] += k
        if with_macro:
            macro[:] += k

    for c1, c2, kind in c08_pairs(g).astype(np.int64):
        i1 = order[start[c1]:start[c1 + 1]]
        if kind == 0:                                                   # processCell
            if not is_halo(c1) and len(i1) >= 2:
                add(i1, i1, True, True)
            continue
        i2 = order[start[c2]:start[c2 + 1]]
        h1, h2 = is_halo(c1), is_halo(c2)
        if h1 and h2:
            continue
        add(i1, i2, False, (not h1 and not h2) or c1 < c2)              # the traversal passes the non-halo cell first
    flops = 8.0 * mm + 8.0 * kernel.sum() + (kernel * (FLOP_MULT[:, 0] + FLOP_MULT[:, 2])).sum() + (macro * (FLOP_MULT[:, 1] + FLOP_MULT[:, 3])).sum()
    return dict(molecule_distances=mm, kernel=kernel, macro=macro, flops=float(flops))


def _rotate(q, d):
    """Quaternion::rotate (Quaternion.cpp:43-62), q = (w, x, y, z) rows, d one body-frame vector"""
    w, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    ww, xx, yy, zz = w * w, x * x, y * y, z * z
    wx, wy, wz, xy, xz, yz = w * x, w * y, w * z, x * y, x * z, y * z
    return np.stack([(ww + xx - yy - zz) * d[0] + 2. * (xy - wz) * d[1] + 2. * (wy + xz) * d[2],
                     2. * (wz + xy) * d[0] + (ww - xx + yy - zz) * d[1] + 2. * (yz - wx) * d[2],
                     2. * (xz - wy) * d[0] + 2. * (wx + yz) * d[1] + (ww - xx - yy + zz) * d[2]], axis=1)


def site_table(sysm):
    """body-frame offsets of all sites in Component::numSites order (LJ, charges, dipoles, quadrupoles; FullMolecule.h:171-186):
    returns (site0[nc + 1], d[S, 3])"""
    rows, site0 = [], [0]
    o = np.zeros(4, dtype=np.int64)
    for c in sysm.counts:
        for t, tab in enumerate((sysm.lj, sysm.ch, sysm.di, sysm.qu)):
            rows.append(tab[o[t]:o[t] + c[t], :3])
            o[t] += c[t]
        site0.append(site0[-1] + int(c[:4].sum()))
    return np.array(site0, dtype=np.int64), (np.concatenate(rows) if rows else np.zeros((0, 3)))


def rdf(sysm, g, cid, r, q, bins, interval, angular_bins=1):
    """numpy restatement of one traversal with the reference's RDFCellProcessor feeding the RDF plugin's counters
    (adapter/RDFCellProcessor.cpp:18-133, io/RDF.h:95-160, MoleculeInterface.h:157-190).  cid, r, q: all molecules incl. halo copies.
    processCell (non-halo cells): pairs i < j; processCellPair(c1, c2) with the non-halo cell first: all pairs, except that a
    (boundary, halo) cell pair only counts when index(c1) < index(c2).  A pair with centre distance^2 dd < rc^2 is observed:
      molecule RDF  bin floor(sqrt(dd)/interval) of component pair (min, max), skipped when dd > (interval*bins)^2;
      site RDF      (when either molecule has more than one site) every site pair (m, n), absolute site positions, same binning,
                    canonical component order; same component and m != n counts [m][n] and [n][m];
      angular RDF   (angular_bins > 1) cos(phi) between each molecule's body z axis and the centre-centre vector.
    Returns dict(mol[nc(nc+1)/2, bins] in (i, j-i) order, site[S, S, bins] dense over global site rows, ardf[nc, nc, bins*angular_bins])."""
    cid = np.asarray(cid, dtype=np.int64)
    r = np.ascontiguousarray(r, dtype=np.float64).reshape(-1, 3)
    q = np.ascontiguousarray(q, dtype=np.float64).reshape(-1, 4)
    nc = len(sysm.counts)
    cell = cell_indices(g, r).astype(np.int64)
    cpd = np.array(list(g.cpd), dtype=np.int64)
    order = np.argsort(cell, kind="stable")
    start = np.searchsorted(cell[order], np.arange(int(cpd.prod()) + 1))
    site0, sd = site_table(sysm)
    S = int(site0[-1])
    nsites = np.diff(site0)
    qn = q / np.sqrt((q * q).sum(axis=1))[:, None]                     # FullMolecule::normalizeQuaternion before the site positions
    pair_row = {}
    for i in range(nc):
        for k in range(nc - i):
            pair_row[(i, i + k)] = len(pair_row)
    mol = np.zeros((len(pair_row), bins), dtype=np.int64)
    site = np.zeros((S, S, bins), dtype=np.int64)
    ardf = np.zeros((nc, nc, bins * angular_bins), dtype=np.int64)
    rc2 = sysm.rc * sysm.rc
    max2 = interval * interval * bins * bins
    aw = 2.0 / angular_bins

    def is_halo(c):
        x, y, z = c % cpd[0], (c // cpd[0]) % cpd[1], c // (cpd[0] * cpd[1])
        return x == 0 or y == 0 or z == 0 or x == cpd[0] - 1 or y == cpd[1] - 1 or z == cpd[2] - 1

    def cos_phi(a, b, dd):                                              # a.orientationAngle(b): dr = r_b - r_a
        qa = q[a]                                                       # (w, x, y, z), not re-normalised (MoleculeInterface.h:167-190)
        ov = np.stack([2. * (qa[:, 1] * qa[:, 3] + qa[:, 0] * qa[:, 2]), 2. * (qa[:, 2] * qa[:, 3] - qa[:, 0] * qa[:, 1]),
                       1. - 2. * (qa[:, 1] * qa[:, 1] + qa[:, 2] * qa[:, 2])], axis=1)
        dr = r[b] - r[a]
        ov2 = ov[:, 0] * ov[:, 0] + ov[:, 1] * ov[:, 1] + ov[:, 2] * ov[:, 2]
        c = np.zeros(len(a))
        for d in range(3):
            c = c + ov[:, d] * dr[:, d] / np.sqrt(ov2) / np.sqrt(dd)
        return c

    def observe(a, b):                                                  # molecule1 = a, molecule2 = b (index arrays of equal length)
        d = r[a] - r[b]                                                 # molecule2.dist2(molecule1): r1 - r2
        dd = d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1] + d[:, 2] * d[:, 2]
        keep = dd < rc2
        a, b, dd = a[keep], b[keep], dd[keep]
        if len(a) == 0:
            return
        inr = ~(dd > max2)
        lo, hi = np.minimum(cid[a], cid[b]), np.maximum(cid[a], cid[b])
        rows = np.array([pair_row[(int(x), int(y))] for x, y in zip(lo[inr], hi[inr])], dtype=np.int64)
        np.add.at(mol, (rows, np.floor(np.sqrt(dd[inr]) / interval).astype(np.int64)), 1)
        if angular_bins > 1 and inr.any():
            a2, b2, d2 = a[inr], b[inr], dd[inr]
            c1, c2 = cos_phi(a2, b2, d2), cos_phi(b2, a2, d2)
            rb = np.floor(np.sqrt(d2) / interval).astype(np.int64)
            np.add.at(ardf, (cid[a2], cid[b2], rb * angular_bins + np.floor((-c1 + 1.) / aw).astype(np.int64)), 1)
            np.add.at(ardf, (cid[b2], cid[a2], rb * angular_bins + np.floor((-c2 + 1.) / aw).astype(np.int64)), 1)
        for ci in range(nc):
            for cj in range(nc):
                if nsites[ci] + nsites[cj] <= 2:
                    continue
                sel = (cid[a] == ci) & (cid[b] == cj)
                if not sel.any():
                    continue
                aa, bb = a[sel], b[sel]
                for m in range(nsites[ci]):
                    pa = r[aa] + _rotate(qn[aa], sd[site0[ci] + m])
                    for n in range(nsites[cj]):
                        pb = r[bb] + _rotate(qn[bb], sd[site0[cj] + n])
                        ds = pa - pb
                        d2 = ds[:, 0] * ds[:, 0] + ds[:, 1] * ds[:, 1] + ds[:, 2] * ds[:, 2]
                        ok = ~(d2 > max2)
                        bn = np.floor(np.sqrt(d2[ok]) / interval).astype(np.int64)
                        ga, gb = (site0[ci] + m, site0[cj] + n) if ci <= cj else (site0[cj] + n, site0[ci] + m)
                        np.add.at(site, (np.full(len(bn), ga), np.full(len(bn), gb), bn), 1)
                        if ci == cj and m != n:
                            np.add.at(site, (np.full(len(bn), gb), np.full(len(bn), ga), bn), 1)

    for c1, c2, kind in c08_pairs(g).astype(np.int64):
        i1 = order[start[c1]:start[c1 + 1]]
        if kind == 0:
            if not is_halo(c1) and len(i1) >= 2:
                ia, ib = np.triu_indices(len(i1), 1)
                observe(i1[ia], i1[ib])
            continue
        i2 = order[start[c2]:start[c2 + 1]]
        h1, h2 = is_halo(c1), is_halo(c2)
        if (h1 and h2) or len(i1) == 0 or len(i2) == 0:
            continue
        if h1:                                                          # the traversal passes the non-halo cell first
            c1, c2, i1, i2, h1, h2 = c2, c1, i2, i1, h2, h1
        if h2 and not c1 < c2:
            continue
        ia, ib = np.meshgrid(np.arange(len(i1)), np.arange(len(i2)), indexing="ij")
        observe(i1[ia.ravel()], i2[ib.ravel()])
    return dict(mol=mol, site=site, ardf=ardf)
