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
