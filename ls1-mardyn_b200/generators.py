"""Synthetic inputs of the benchmark configurations: 1CLJ fcc lattice, rho*=0.8442, as the reference's
generators build it (utils/generator/Lattice.h:31-61 face-centred cubic; velocity rule ParticleCellBase.cpp:33-53:
direction uniform in [-1/2,1/2]^3, |v|^2 = 3T).  Everything is a pure function of the GLOBAL molecule id, so any
block decomposition generates exactly the same system (ids, positions, velocities) without communication."""
import numpy as np

_FCC = np.array([[0.0, 0.0, 0.0], [0.5, 0.5, 0.0], [0.5, 0.0, 0.5], [0.0, 0.5, 0.5]])


def fcc_box(n_cells, rho=0.8442):
    """(lattice constant, box length[3]) of an fcc lattice with n_cells unit cells per dimension"""
    a = (4.0 / rho) ** (1.0 / 3.0)
    n = np.asarray(n_cells, dtype=np.int64) * np.ones(3, dtype=np.int64)
    return a, n.astype(np.float64) * a


def _splitmix64(x):
    x = (x + np.uint64(0x9E3779B97F4A7C15))
    z = x
    z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return z ^ (z >> np.uint64(31))


def velocities_of_ids(ids, temperature, seed=1):
    """deterministic per-molecule velocity from the global id (counter-based hash)"""
    ids = np.asarray(ids, dtype=np.uint64)
    with np.errstate(over="ignore"):
        v = np.empty((len(ids), 3))
        for d in range(3):
            h = _splitmix64(ids * np.uint64(3) + np.uint64(d) + np.uint64(seed) * np.uint64(0x100000001B3))
            v[:, d] = (h >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0) - 0.5
    n2 = np.maximum((v * v).sum(axis=1), 1e-300)
    v *= np.sqrt(3.0 * temperature / n2)[:, None]
    return v


def fcc_block(n_cells, box_min=None, box_max=None, rho=0.8442, temperature=1.44, seed=1):
    """molecules of the global fcc lattice (n_cells^3 unit cells) whose position lies in [box_min, box_max).
    returns (id uint64, r[n][3], v[n][3]); id = 4*((i*ny + j)*nz + k) + basis index."""
    a, L = fcc_box(n_cells, rho)
    n = np.asarray(n_cells, dtype=np.int64) * np.ones(3, dtype=np.int64)
    lo = np.zeros(3) if box_min is None else np.asarray(box_min, dtype=np.float64)
    hi = L if box_max is None else np.asarray(box_max, dtype=np.float64)
    rng_idx = []
    for d in range(3):
        i0 = max(int(np.floor(lo[d] / a)) - 1, 0)
        i1 = min(int(np.ceil(hi[d] / a)) + 1, int(n[d]))
        rng_idx.append(np.arange(i0, i1, dtype=np.int64))
    ids_out, r_out = [], []
    # slab-wise over x to bound temporary memory at 10^7..10^8 molecules
    for i in rng_idx[0]:
        j, k = np.meshgrid(rng_idx[1], rng_idx[2], indexing="ij")
        j, k = j.ravel(), k.ravel()
        base = np.stack([np.full_like(j, i), j, k], axis=1).astype(np.float64)
        cell_id = (i * n[1] + j) * n[2] + k
        r = ((base[:, None, :] + _FCC[None, :, :]) * a + 0.25 * a).reshape(-1, 3)
        ids = (cell_id[:, None] * 4 + np.arange(4)[None, :]).reshape(-1)
        keep = np.all((r >= lo) & (r < hi), axis=1)
        ids_out.append(ids[keep].astype(np.uint64))
        r_out.append(r[keep])
    ids = np.concatenate(ids_out) if ids_out else np.zeros(0, dtype=np.uint64)
    r = np.concatenate(r_out) if r_out else np.zeros((0, 3))
    return ids, r, velocities_of_ids(ids, temperature, seed)
