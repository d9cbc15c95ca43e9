"""Binary checkpoint written by the REFERENCE itself (test infrastructure; needs oracle/_ref/MarDyn, i.e. the build container).

Runs the unmodified reference executable on the per_MultiComponentMultiPotentials fixture (250 molecules, 5 components, rigid
rotation) for one time step with its CheckpointWriter plugin in binary mode (io/CheckpointWriter.cpp:28-33 ->
Domain::writeCheckpoint -> DomainDecompBase::writeMoleculesToFile -> FullMolecule::writeBinary) and keeps
    tests/golden/cp_binary_MultiComponent.dat / .header.xml   the binary checkpoint after that step, byte for byte
    tests/golden/cp_ascii_MultiComponent.restart.dat          the same state as the reference's ASCII writer printed it (final
                                                              checkpoint, FullMolecule::write at 20 digits), molecules in the same order
    tests/golden/cp_binary_MultiComponent.npz                 ... parsed: id cid r v q D
so the two writers of the reference pin each other (the script checks that they agree exactly; the checkpoint of step 0 is
checked against the fixture's input molecules).  The record layout of ls1g_upload_records / ls1g_download_records (and of phasespace.read_binary / write_binary) is pinned to
these bytes: tests/test_records.py.

    python tests/golden/make_golden_r3.py
"""
import os
import shutil
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from tests import mardyn_cases as M          # noqa: E402
from tests.helpers import load               # noqa: E402

NAME = "per_MultiComponentMultiPotentials"
PLUGIN = ('<outputplugin name="CheckpointWriter"><type>binary</type><writefrequency>1</writefrequency>'
          '<outputprefix>cpbin</outputprefix></outputplugin>')


def main():
    out = os.path.join(ROOT, "tests", "golden")
    with tempfile.TemporaryDirectory() as tmp:
        cfg = M.write_case(tmp, NAME, 1, "cpu", temperature=0.0007, output_plugins=PLUGIN)
        # the fixture's time step (0.5, chosen for its force comparison) throws half of the molecules out of the halo in one step
        xml = open(cfg).read()
        assert '<timestep unit="reduced">0.5</timestep>' in xml
        open(cfg, "w").write(xml.replace('<timestep unit="reduced">0.5</timestep>', '<timestep unit="reduced">0.01</timestep>'))
        res = M.run(M.REF_EXE, cfg, 1, threads=1)
        dtype = np.dtype([("id", "<u8"), ("cid", "<u4"), ("r", "<f8", 3), ("v", "<f8", 3), ("q", "<f8", 4), ("D", "<f8", 3)])
        # step 0: what the reference's ASCIIReader took from the fixture's phase-space file
        d = load(NAME)
        rec = np.fromfile(os.path.join(tmp, "cpbin-0.restart.dat"), dtype=dtype)
        o, oi = np.argsort(rec["id"]), np.argsort(d["in_id"])
        assert len(rec) == len(d["in_id"]) and np.array_equal(rec["id"][o], d["in_id"][oi].astype(np.uint64))
        assert np.array_equal(rec["cid"][o], d["in_cid"][oi] + 1)
        assert np.array_equal(rec["r"][o], d["in_r"][oi]) and np.array_equal(rec["v"][o], d["in_v"][oi])
        # step 1: the binary writer against the ASCII writer
        dat, hdr = os.path.join(tmp, "cpbin-1.restart.dat"), os.path.join(tmp, "cpbin-1.restart.header.xml")
        rec = np.fromfile(dat, dtype=dtype)
        o, f = np.argsort(rec["id"]), res["final"]
        assert len(rec) == len(f["id"]) and np.array_equal(rec["id"][o], f["id"]) and np.array_equal(rec["cid"][o], f["cid"] + 1)
        for a in ("r", "v", "q", "D"):
            assert np.array_equal(rec[a][o], f[a]), a
        np.savez_compressed(os.path.join(out, "cp_binary_MultiComponent.npz"), **f)
        # the ASCII checkpoint itself (header + one FullMolecule::write line per molecule, stream precision 20), same state, same order
        asc = os.path.join(tmp, NAME + "_cpu.restart.dat")
        lines = open(asc).read().split("\n")
        k = next(i for i, l in enumerate(lines) if l.split()[:1] == ["MoleculeFormat"])
        assert [int(l.split()[0]) for l in lines[k + 1:] if l] == rec["id"].tolist()
        shutil.copy(asc, os.path.join(out, "cp_ascii_MultiComponent.restart.dat"))
        shutil.copy(dat, os.path.join(out, "cp_binary_MultiComponent.dat"))
        shutil.copy(hdr, os.path.join(out, "cp_binary_MultiComponent.header.xml"))
    print("wrote cp_binary_MultiComponent.{dat,header.xml}: %d records" % len(rec))


if __name__ == "__main__":
    main()
