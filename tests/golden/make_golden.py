"""Generate the golden fixtures under tests/golden/ by running the UNMODIFIED reference
(oracle/_ref/ref_driver, built by oracle/ref/Makefile from /root/reference/src).

Run in the build container only (needs /root/reference):   python tests/golden/make_golden.py
The GPU box never reads /root/reference; it uses the committed .npz files.

Cases
  vcp_<name>     the reference's own VectorizedCellProcessorTest inputs (test_input/Vectorization*.inp,
                 ForceCalculationTest{U0,F0}.inp), set up exactly like its test factory: LinkedCells
                 constructor, no halo copies ("open"), rc as in VectorizedCellProcessorTest.cpp:349-387.
  per_<name>     the same physics with periodic images (DomainDecompBase::exchangeMolecules), XML-path
                 grid (LinkedCells::rebuild), plus N leapfrog steps NVE.
  ex_<name>      shipped examples (Argon 1CLJ, CO2 3LJ+3q, EOX 3LJ+dipole), periodic, rebuild grid, 5 steps.
  cellindex_*    adversarial points on/near cell borders -> LinkedCells::getCellIndexOfPoint.
  rdf_<name>     RDFCellProcessor + RDF plugin counters of one traversal (molecule pairs per component pair, site-site, angular),
                 driven like the reference's RDFTest (io/tests/RDFTest.cpp: 1clj-regular-2x2x3, 1clj-regular-12x12x12, 2clj-regular
                 with its expected bin counts) and on the shipped examples / multi-component inputs.   --rdf-only writes just these.
"""
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("LS1_REFERENCE", "/root/reference")
DRIVER = os.path.join(ROOT, "oracle", "_ref", "ref_driver")
sys.path.insert(0, HERE)
from refdump import read_dump  # noqa: E402

TI = os.path.join(REF, "test_input")
EX = os.path.join(REF, "examples")

VCP = [  # (name, file, rc)  -- VectorizedCellProcessorTest.cpp:45-134,349-387
    ("U0", "ForceCalculationTestU0.inp", 1.1),
    ("F0", "ForceCalculationTestF0.inp", 1.3),
    ("LennardJones", "VectorizationLennardJones.inp", 35.0),
    ("Charge", "VectorizationCharge.inp", 35.0),
    ("ChargeDipole", "VectorizationChargeDipole.inp", 35.0),
    ("ChargeQuadrupole", "VectorizationChargeQuadrupole.inp", 35.0),
    ("Dipole", "VectorizationDipole.inp", 35.0),
    ("DipoleQuadrupole", "VectorizationDipoleQuadrupole.inp", 35.0),
    ("Quadrupole", "VectorizationQuadrupole.inp", 35.0),
    ("Water", "VectorizationWater.inp", 6.16),
    ("MultiComponentMultiPotentials", "VectorizationMultiComponentMultiPotentials.inp", 35.0),
]
PER = [("LennardJones", "VectorizationLennardJones.inp", 35.0, 3, 0.5),
       # components with zero mass (pure multipole test molecules) cannot be time-stepped: forces only
       ("MultiComponentMultiPotentials", "VectorizationMultiComponentMultiPotentials.inp", 35.0, 0, 0.5),
       ("Water", "VectorizationWater.inp", 6.16, 3, 0.002),
       ("Dipole", "VectorizationDipole.inp", 35.0, 0, 0.5)]
EXAMPLES = [("Argon", "Argon/200K_18mol_l/Argon_200K_18mol_l.inp", 33.0702, 5, 0.0667516),
            ("CO2", "CO2/370K_15mol_l/CO2_370K_15mol_l.inp", 33.0702, 5, 0.0667516),
            ("EOX", "EOX/600K_15mol_l/EOX_600K_15mol_l.inp", 33.0702, 5, 0.0667516)]


def run_driver(args, threads=2):
    env = dict(os.environ, OMP_NUM_THREADS=str(threads))
    subprocess.check_call([DRIVER] + [str(a) for a in args], env=env)


def is_halo(cell, cpd):
    x = cell % cpd[0]
    y = (cell // cpd[0]) % cpd[1]
    z = cell // (cpd[0] * cpd[1])
    return (x == 0) | (y == 0) | (z == 0) | (x == cpd[0] - 1) | (y == cpd[1] - 1) | (z == cpd[2] - 1)


def case(name, inp, rc, periodic, how, steps, dt):
    with tempfile.TemporaryDirectory() as td:
        out = os.path.join(td, "o.bin")
        run_driver(["run", inp, rc, rc, "periodic" if periodic else "open", how, steps, dt, out])
        d = read_dump(out)
    cpd = d["cells_per_dim"]
    keep = {k: d[k] for k in ("comp_counts", "comp_scalars", "lj_sites", "charge_sites", "dipole_sites",
                              "quadrupole_sites", "mixcoeff", "epsRF", "box", "rc", "dt", "cells_per_dim",
                              "cell_length", "upot_virial", "flops")}
    for k in ("id", "cid", "r", "v", "q", "L"):
        keep["in_" + k] = d["in_" + k]
    keep["in_cell"] = d["in_cell"]
    # force pass: all molecules incl. halo copies, in container order (owned flagged)
    halo = is_halo(d["f_cell"], cpd)
    keep["f_owned"] = (~halo).astype(np.uint8)
    for k in ("id", "cid", "cell", "r", "q", "F", "M", "Vi"):
        keep["f_" + k] = d["f_" + k]
    # cell-pair multiset, sorted (order inside a colour is schedule dependent in the reference)
    cp = d["cellpairs"]
    keep["cellpairs"] = cp[np.lexsort((cp[:, 2], cp[:, 1], cp[:, 0]))].astype(np.uint32)
    if steps > 0:
        keep["series"] = d["series"]
        for k in ("id", "r", "v", "q", "L", "F", "M"):
            keep["post_" + k] = d["post_" + k]
    keep["meta"] = np.array([int(periodic), 1 if how == "ctor" else 0, steps], dtype=np.int64)
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **keep)
    print(name, os.path.getsize(path) // 1024, "KiB")


def cellindex_case(name, L, rc, how, seed):
    rng = np.random.default_rng(seed)
    # borders as the reference computes them are unknown here; sample around k*cell for both
    # candidate cell counts plus random points and points a few ulps around each candidate border
    pts = [rng.uniform(-0.3 * rc, np.array(L) + 0.3 * rc, size=(2000, 3))]
    for d in range(3):
        n = int(np.floor(L[d] / (np.float32(rc) if how == "rebuild" else rc)))
        cl = L[d] / n
        for k in range(-1, n + 2):
            for b in (k * cl, k * cl - cl + cl, (k + 1) * cl - cl, np.float64(k) * cl + (-cl) + cl):
                for u in range(-3, 4):
                    x = b
                    for _ in range(abs(u)):
                        x = np.nextafter(x, np.inf if u > 0 else -np.inf)
                    p = rng.uniform(0, L)
                    p[d] = x
                    pts.append(p[None, :])
    pts = np.ascontiguousarray(np.concatenate(pts))
    with tempfile.TemporaryDirectory() as td:
        pf, of = os.path.join(td, "p.bin"), os.path.join(td, "o.bin")
        pts.tofile(pf)
        run_driver(["cellindex", L[0], L[1], L[2], rc, how, pf, of])
        d = read_dump(of)
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, points=pts, box=np.array(L, dtype=np.float64), rc=np.array([rc]),
                        mode=np.array([1 if how == "ctor" else 0]), cells_per_dim=d["cells_per_dim"],
                        cell_length=d["cell_length"], cell_index=d["cell_index"].astype(np.uint32))
    print(name, os.path.getsize(path) // 1024, "KiB", "cpd", d["cells_per_dim"])


RDF = [  # (name, file, rc, periodic, how, bins, intervallength, angular bins)
    ("1clj_2x2x3_open", os.path.join(TI, "1clj-regular-2x2x3.inp"), 1.8, False, "ctor", 100, 0.018, 1),      # RDFTest.cpp:52-90
    ("1clj_2x2x3", os.path.join(TI, "1clj-regular-2x2x3.inp"), 1.8, True, "ctor", 100, 0.018, 1),            # RDFTest.cpp:92-120
    ("1clj_12x12x12", os.path.join(TI, "1clj-regular-12x12x12.inp"), 1.8, True, "ctor", 100, 0.018, 1),      # RDFTest.cpp:124-160
    ("2clj", os.path.join(TI, "2clj-regular.inp"), 3.5, True, "ctor", 101, 0.05, 1),                         # RDFTest.cpp:205-260
    ("ex_Argon", os.path.join(EX, "Argon/200K_18mol_l/Argon_200K_18mol_l.inp"), 33.0702, True, "rebuild", 64, 0.6, 1),
    ("ex_CO2", os.path.join(EX, "CO2/370K_15mol_l/CO2_370K_15mol_l.inp"), 33.0702, True, "rebuild", 50, 0.5, 8),
    ("ex_EOX", os.path.join(EX, "EOX/600K_15mol_l/EOX_600K_15mol_l.inp"), 33.0702, True, "rebuild", 40, 0.9, 5),
    ("Multi", os.path.join(TI, "VectorizationMultiComponentMultiPotentials.inp"), 35.0, True, "rebuild", 30, 1.0, 4),
    ("DipoleQuadrupole", os.path.join(TI, "VectorizationDipoleQuadrupole.inp"), 35.0, False, "ctor", 25, 1.5, 6),
]


def rdf_case(name, inp, rc, periodic, how, bins, interval, abins):
    with tempfile.TemporaryDirectory() as td:
        out = os.path.join(td, "o.bin")
        run_driver(["rdf", inp, rc, "periodic" if periodic else "open", how, bins, interval, abins, out])
        d = read_dump(out)
    keep = {k: d[k] for k in ("comp_counts", "comp_scalars", "lj_sites", "charge_sites", "dipole_sites", "quadrupole_sites",
                              "mixcoeff", "epsRF", "box", "rc", "cells_per_dim", "rdf_params")}
    halo = is_halo(d["pre_cell"], d["cells_per_dim"])
    keep["f_owned"] = (~halo).astype(np.uint8)
    for k in ("id", "cid", "cell", "r", "q"):        # all molecules incl. halo copies, container order (as the f_ rows of the force cases)
        keep["f_" + k] = d["pre_" + k]
    keep["rdf_mol"] = d["rdf_mol"].astype(np.uint32)
    keep["rdf_site"] = d["rdf_site"].astype(np.uint32)
    keep["rdf_ardf"] = d["rdf_ardf"].astype(np.uint32)
    keep["meta"] = np.array([int(periodic), 1 if how == "ctor" else 0, 0], dtype=np.int64)
    path = os.path.join(HERE, "rdf_" + name + ".npz")
    np.savez_compressed(path, **keep)
    print("rdf_" + name, os.path.getsize(path) // 1024, "KiB", "pairs", int(keep["rdf_mol"].sum()), "site", int(keep["rdf_site"].sum()),
          "ardf", int(keep["rdf_ardf"].sum()))


def main():
    if "--rdf-only" in sys.argv:
        for c in RDF:
            rdf_case(*c)
        return
    for name, f, rc in VCP:
        case("vcp_" + name, os.path.join(TI, f), rc, False, "ctor", 0, 0.001)
    for name, f, rc, steps, dt in PER:
        case("per_" + name, os.path.join(TI, f), rc, True, "rebuild", steps, dt)
    for name, f, rc, steps, dt in EXAMPLES:
        case("ex_" + name, os.path.join(EX, f), rc, True, "rebuild", steps, dt)
    cellindex_case("cellindex_rebuild_a", [108.43455, 108.43455, 108.43455], 33.0702, "rebuild", 1)
    cellindex_case("cellindex_rebuild_b", [22.8421, 17.3, 30.0], 2.5, "rebuild", 2)
    cellindex_case("cellindex_ctor_a", [10.0, 10.0, 10.0], 1.0, "ctor", 3)
    cellindex_case("cellindex_ctor_b", [7.7, 9.1, 13.3], 1.1, "ctor", 4)
    for c in RDF:
        rdf_case(*c)


if __name__ == "__main__":
    main()
