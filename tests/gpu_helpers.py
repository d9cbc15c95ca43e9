"""Helpers of the GPU parity tests: build a product Container (C ABI, CUDA) from a golden dump or a seeded
synthetic system.  Test infrastructure."""
import importlib

import numpy as np

ls1g = importlib.import_module("ls1-mardyn_b200")


def components_of(d):
    return ls1g.components_from_tables(d["comp_counts"], d["comp_scalars"], d["lj_sites"], d["charge_sites"],
                                       d["dipole_sites"], d["quadrupole_sites"])


def container_of(d, **kw):
    """Container with the grid, cutoffs and component tables of a golden dump"""
    c = ls1g.Container(d["box"], float(d["rc"][0]), lj_cutoff=float(d["rc"][1]), eps_rf=float(d["epsRF"][0]),
                       grid_mode=int(d["meta"][1]), **kw)
    c.set_components(components_of(d), d["mixcoeff"])
    return c


def by_id(out):
    """sort a download() dict by molecule id"""
    order = np.argsort(out["id"], kind="stable")
    return {k: v[order] for k, v in out.items()}


def fcc_lattice(n_cells, rho=0.8442, temperature=1.44, seed=1):
    """cfg B/E style input: fcc lattice, a=(4/rho)^(1/3), positions (i+{0,1/2},...)*a + a/4;
    velocities uniform in [-1/2,1/2]^3 rescaled to |v|^2 = 3T (ParticleCellBase.cpp:33-53), net momentum removed."""
    a = (4.0 / rho) ** (1.0 / 3.0)
    n = np.asarray(n_cells, dtype=np.int64) * np.ones(3, dtype=np.int64)
    i, j, k = np.meshgrid(np.arange(n[0]), np.arange(n[1]), np.arange(n[2]), indexing="ij")
    base = np.stack([i.ravel(), j.ravel(), k.ravel()], axis=1).astype(np.float64)
    offs = np.array([[0, 0, 0], [0.5, 0.5, 0], [0.5, 0, 0.5], [0, 0.5, 0.5]])
    r = (base[:, None, :] + offs[None, :, :]).reshape(-1, 3) * a + 0.25 * a
    rng = np.random.default_rng(seed)
    v = rng.uniform(-0.5, 0.5, size=r.shape)
    v *= np.sqrt(3.0 * temperature / (v * v).sum(axis=1))[:, None]
    v -= v.mean(axis=0)
    box = n.astype(np.float64) * a
    return r, v, box


def random_quaternions(n, rng):
    q = rng.normal(size=(n, 4))
    return q / np.linalg.norm(q, axis=1)[:, None]
