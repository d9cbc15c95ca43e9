"""the reference's binary phase-space records on the device boundary (ls1g_upload_records / ls1g_download_records,
ls1-mardyn_b200/csrc/records.cu) and the Python reader / writer of the same format (phasespace.read_binary / write_binary).

Pinned to a checkpoint the REFERENCE wrote (tests/golden/cp_binary_MultiComponent.dat + .header.xml: CheckpointWriter, binary
mode, FullMolecule::writeBinary; tests/golden/make_golden_r3.py) and to the same state as the reference's ASCII writer printed
it (.npz).  Bar: byte / bit exact - records are copied, not computed.  The GPU tests go through the C ABI and compare with the
reference's bytes and with an independent numpy decoding of the packed layout, never with the library itself."""
import os

import numpy as np
import pytest

from tests.helpers import GOLDEN, load
from tests.gpu_helpers import ls1g, container_of, fcc_lattice, random_quaternions

P = ls1g.phasespace
DAT = os.path.join(GOLDEN, "cp_binary_MultiComponent.dat")
HDR = os.path.join(GOLDEN, "cp_binary_MultiComponent.header.xml")
FORMATS = {"ICRVQD": P.BINARY_RECORD,
           "ICRV": np.dtype([("id", "<u8"), ("cid", "<u4"), ("r", "<f8", 3), ("v", "<f8", 3)]),
           "IRV": np.dtype([("id", "<u8"), ("r", "<f8", 3), ("v", "<f8", 3)])}


def reference_records():
    return np.fromfile(DAT, dtype=P.BINARY_RECORD)


# ---------------------------------------------------------------------------------------------------------------------------
# CPU: the layout itself, and the host-only entry point
# ---------------------------------------------------------------------------------------------------------------------------
def test_binary_and_ascii_writer_of_the_reference_agree():
    """the fixture pins itself: 116-byte records, and the fields decoded from them equal what the reference's ASCII writer printed"""
    rec, asc = reference_records(), load("cp_binary_MultiComponent")
    assert os.path.getsize(DAT) == 116 * len(rec) == 116 * len(asc["id"]) and len(rec) == 128
    o = np.argsort(rec["id"])
    assert np.array_equal(rec["id"][o], asc["id"]) and np.array_equal(rec["cid"][o], asc["cid"] + 1)
    for k in ("r", "v", "q", "D"):
        assert np.array_equal(rec[k][o], asc[k]), k
    assert len(np.unique(rec["cid"])) == 3 and np.abs(rec["D"]).max() > 0          # several components, rotating molecules


def test_python_reader_and_writer_reproduce_the_reference_files(tmp_path):
    ps = P.read_binary(HDR, DAT, n_components=5)
    rec = reference_records()
    assert ps["time"] == 0.01 and np.array_equal(ps["box"], load("per_MultiComponentMultiPotentials")["box"])
    assert np.array_equal(ps["id"], rec["id"]) and np.array_equal(ps["cid"] + 1, rec["cid"])
    for k in ("r", "v", "q", "D"):
        assert np.array_equal(ps[k], rec[k])
    P.write_binary(tmp_path / "w.header.xml", tmp_path / "w.dat", ps)
    assert open(tmp_path / "w.dat", "rb").read() == open(DAT, "rb").read()
    assert open(tmp_path / "w.header.xml", "rb").read() == open(HDR, "rb").read()      # Domain::writeCheckpointHeaderXML, byte for byte
    with pytest.raises(ValueError):
        P.read_binary(HDR, DAT, n_components=4)                                         # BinaryReader.cpp:222-228


def test_ascii_lines_printed_from_the_packed_records_are_the_reference_writers(tmp_path):
    """host/RecordText.h (what GpuLinkedCells::writeRecords prints for <type>ASCII</type>): the molecule lines of the reference's
    ASCII checkpoint, byte for byte, from the reference's binary checkpoint of the same state"""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = str(tmp_path / "record_text_check")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-I", os.path.join(root, "ls1-mardyn_b200", "host"),
                           os.path.join(root, "tests", "record_text_check.cpp"), "-o", exe])
    mine = subprocess.run([exe, DAT], capture_output=True, check=True).stdout
    ref = open(os.path.join(GOLDEN, "cp_ascii_MultiComponent.restart.dat"), "rb").read()
    k = ref.index(b"MoleculeFormat")
    body = ref[ref.index(b"\n", k) + 1:]
    assert mine == body and mine.count(b"\n") == 128


@pytest.mark.parametrize("n_ranks", [2, 4, 8, 6])
def test_bounding_boxes_of_the_ranks_partition_a_file(n_ranks):
    """ls1g_upload_records keeps min <= r < max of the rank's box (isInBoundingBox): with DomainDecomposition's boxes (ls1g_decomp_grid /
    ls1g_decomp_box, host only) every record of a file lands on exactly one rank - also records exactly on the faces between ranks"""
    rng = np.random.default_rng(n_ranks)
    box = np.array([31.7, 25.3, 40.9])
    r = rng.uniform(0.0, 1.0, size=(20000, 3)) * box
    owners = np.zeros(len(r), dtype=np.int64)
    faces = []
    for rank in range(n_ranks):
        grid, coords = ls1g.decomp_grid(n_ranks, rank)
        bmin, bmax = ls1g.decomp_box(grid, coords, box)
        faces.append(bmax)
    r[:len(faces)] = np.minimum(np.array(faces), np.nextafter(box, 0.0))          # points ON the upper faces of the rank boxes
    for rank in range(n_ranks):
        grid, coords = ls1g.decomp_grid(n_ranks, rank)
        bmin, bmax = ls1g.decomp_box(grid, coords, box)
        owners += np.all((r >= bmin) & (r < bmax), axis=1)
    assert np.all(owners == 1)


def test_record_size_is_host_only():
    assert [ls1g.Container.record_size(f) for f in ("ICRVQD", "ICRV", "IRV")] == [116, 60, 56] == [FORMATS[f].itemsize for f in ("ICRVQD", "ICRV", "IRV")]
    with pytest.raises(ls1g.Ls1gError, match="Not a valid molecule format"):
        ls1g.Container.record_size(3)


# ---------------------------------------------------------------------------------------------------------------------------
# GPU, through the C ABI
# ---------------------------------------------------------------------------------------------------------------------------
@pytest.mark.gpu
def test_upload_records_of_the_reference_checkpoint():
    """the reference's .dat bytes go to the device as they are; what comes back field by field equals the reference's ASCII dump"""
    rec, asc = reference_records(), load("cp_binary_MultiComponent")
    with container_of(load("per_MultiComponentMultiPotentials")) as c:
        kept, max_id = c.add_records(open(DAT, "rb").read())
        assert (kept, max_id) == (128, int(rec["id"].max())) and c.count() == (128, 0)
        out = c.download(("id", "cid", "r", "v", "q", "L"))
        assert np.array_equal(out["id"], rec["id"])                                    # file order is kept
        o = np.argsort(out["id"])
        assert np.array_equal(out["cid"][o], asc["cid"])
        for k, a in (("r", "r"), ("v", "v"), ("q", "q"), ("L", "D")):
            assert np.array_equal(out[k][o], asc[a]), k


@pytest.mark.gpu
def test_download_records_are_the_bytes_the_reference_wrote():
    """upload, re-sort into cells with halo copies in between, write records: the reference's checkpoint up to the order of the
    molecules (cell by cell, as the reference's own writer goes through its cells), byte for byte"""
    raw = np.fromfile(DAT, dtype=np.uint8).reshape(-1, 116)
    ids = reference_records()["id"]
    with container_of(load("per_MultiComponentMultiPotentials")) as c:
        c.add_records(raw)
        unsorted = c.download_records()
        assert np.array_equal(unsorted, raw)                                           # before any sort: the file itself
        c.exchange()
        assert c.count()[1] > 0                                                        # halo copies sit between the owned molecules
        got = c.download_records()
        order = c.download(("id",))["id"]
        assert got.shape == (128, 116) and np.array_equal(got.view(P.BINARY_RECORD)["id"].ravel(), order)
        assert np.array_equal(got[np.argsort(order)], raw[np.argsort(ids)])
        c.delete_outer_particles()
        assert np.array_equal(c.download_records(), got)
        # a velocity scaling that the device has not applied yet goes into the records (Molecule::scale_v / scale_D: one
        # multiplication per component), the device state stays as it is
        scaled = c.download_records(scale_v=0.97, scale_D=1.03).view(P.BINARY_RECORD).ravel()
        plain = got.view(P.BINARY_RECORD).ravel()
        assert np.array_equal(scaled["v"], plain["v"] * 0.97) and np.array_equal(scaled["D"], plain["D"] * 1.03)
        for k in ("id", "cid", "r", "q"):
            assert np.array_equal(scaled[k], plain[k])
        assert np.array_equal(c.download_records(), got)


def synthetic_records(n, fmt, n_comp, box, rng, outside=0.03):
    rec = np.zeros(n, dtype=FORMATS[fmt])
    rec["id"] = rng.permutation(n).astype(np.uint64) + 7
    rec["r"] = rng.uniform(0.0, 1.0, size=(n, 3)) * box
    k = int(outside * n)
    if k:
        sel = rng.choice(n, size=k, replace=False)
        rec["r"][sel[: k // 3], 0] = box[0]                                            # exactly on the upper face: outside (min <= r < max)
        rec["r"][sel[k // 3: 2 * k // 3], 1] = -1e-9
        rec["r"][sel[2 * k // 3:], 2] += box[2]
        rec["r"][sel[-1]] = 0.0                                                        # exactly on the lower corner: inside
        rec["id"][sel[0]] = 2 ** 40 + 5                                                # the highest id belongs to a record outside the box
    rec["v"] = rng.normal(size=(n, 3))
    if "cid" in rec.dtype.names:
        rec["cid"] = rng.integers(1, n_comp + 1, size=n)
    if fmt == "ICRVQD":
        rec["q"] = random_quaternions(n, rng)
        rec["D"] = rng.normal(size=(n, 3))
    return rec


@pytest.mark.gpu
@pytest.mark.parametrize("fmt,rot,n", [("ICRVQD", True, 600_001), ("ICRVQD", False, 262_144 + 300), ("ICRV", True, 70_000), ("IRV", False, 262_144),
                                       ("ICRVQD", True, 255), ("ICRV", False, 1)])
def test_upload_records_against_numpy_decoding(fmt, rot, n):
    """several chunks of the staging area (262 144 records each), ragged tails, all three formats, with and without orientations on
    the device, records outside the bounding box dropped in place (BinaryReader.cpp:271-273), max id over ALL records (:281-282)"""
    rng = np.random.default_rng(n)
    box = np.full(3, 40 * (4.0 / 0.8442) ** (1.0 / 3.0))          # the box of a 40^3 fcc lattice: 27^3 cells of 2.5 sigma
    comps = [ls1g.lj_component()]              # one single-centre component: the device keeps no orientations
    if rot:
        comps.append(dict(ls1g.lj_component(sigma=1.1), inertia=[1.0, 2.0, 3.0], rot_dof=3,
                          lj=np.array([[0.3, 0, 0, 1.0, 1.0, 1.0, 0.0], [-0.3, 0, 0, 1.0, 1.0, 1.0, 0.0]])))
    n_comp = 1 if fmt == "IRV" else len(comps)
    rec = synthetic_records(n, fmt, n_comp, box, rng, outside=0.03 if n > 100 else 0.0)
    with ls1g.Container(box, 2.5) as c:
        c.set_components(comps, [1.0, 1.0] if rot else ())
        kept, max_id = c.add_records(rec, fmt)
        inside = np.all((rec["r"] >= 0.0) & (rec["r"] < box), axis=1)
        assert kept == int(inside.sum()) and max_id == int(rec["id"].max())
        assert n <= 100 or (kept < n and not inside[np.argmax(rec["id"])])
        out = c.download(("id", "cid", "r", "v", "q", "L"))
        want = rec[inside]
        assert np.array_equal(out["id"], want["id"]) and np.array_equal(out["r"], want["r"]) and np.array_equal(out["v"], want["v"])
        assert np.array_equal(out["cid"] + 1, want["cid"] if "cid" in rec.dtype.names else np.ones(kept, dtype=np.uint32))
        unit = np.tile([1.0, 0.0, 0.0, 0.0], (kept, 1))
        full = fmt == "ICRVQD" and rot
        assert np.array_equal(out["q"], want["q"] if full else unit) and np.array_equal(out["L"], want["D"] if full else np.zeros((kept, 3)))
        # and out again as ICRVQD records, before and after a sort (file order / cell order)
        back = c.download_records().view(P.BINARY_RECORD).ravel()
        expect = np.zeros(kept, dtype=P.BINARY_RECORD)
        expect["id"], expect["cid"], expect["r"], expect["v"] = out["id"], out["cid"] + 1, out["r"], out["v"]
        expect["q"], expect["D"] = out["q"], out["L"]
        assert back.tobytes() == expect.tobytes()
        if n > 100_000:
            c.exchange()
            sorted_records = c.download_records().view(P.BINARY_RECORD).ravel()
            order = c.download(("id",))["id"]
            assert np.array_equal(sorted_records["id"], order)
            assert sorted_records[np.argsort(order)].tobytes() == expect[np.argsort(expect["id"])].tobytes()


@pytest.mark.gpu
def test_upload_records_equals_upload_of_the_arrays():
    """with every record inside the box the context is in the state ls1g_upload leaves: same forces from the same molecules"""
    d = load("ex_EOX")
    rec = np.zeros(len(d["in_id"]), dtype=P.BINARY_RECORD)
    rec["id"], rec["cid"], rec["r"], rec["v"], rec["q"], rec["D"] = d["in_id"], d["in_cid"] + 1, d["in_r"], d["in_v"], d["in_q"], d["in_L"]
    res = []
    for records in (False, True):
        with container_of(d) as c:
            if records:
                assert c.add_records(rec)[0] == len(rec)
            else:
                c.add_particles(d["in_id"], d["in_cid"], d["in_r"], d["in_v"], d["in_q"], d["in_L"])
            c.exchange()
            u, w = c.traverse_cells()
            out = c.download(("id", "F", "M"))
            o = np.argsort(out["id"])
            res.append((u, w, out["F"][o], out["M"][o]))
    assert res[0][0] == res[1][0] and res[0][1] == res[1][1]
    assert np.array_equal(res[0][2], res[1][2]) and np.array_equal(res[0][3], res[1][3])
    own = d["f_owned"].astype(bool)
    ref_F = d["f_F"][own][np.argsort(d["f_id"][own])]
    assert np.abs(res[1][2] - ref_F).max() <= 1e-10 * np.abs(ref_F).max()              # and both are the reference's forces


@pytest.mark.gpu
def test_every_rank_keeps_its_part_of_the_file():
    """BinaryReader hands every rank the whole file; a rank keeps what lies in its bounding box.  Two ranks: a partition of the file."""
    rng = np.random.default_rng(11)
    _, _, box = fcc_lattice(6)
    rec = synthetic_records(20_000, "ICRV", 1, box, rng, outside=0.0)
    parts = []
    for rank in range(2):
        grid, coords = ls1g.decomp_grid(2, rank)
        bmin, bmax = ls1g.decomp_box(grid, coords, box)
        with ls1g.Container(bmax, 2.5, box_min=bmin, global_length=box, rank_grid=grid, rank_coords=coords) as c:
            c.set_components([ls1g.lj_component()])
            kept, max_id = c.add_records(rec, "ICRV")
            ids = c.download(("id",))["id"]
            inside = np.all((rec["r"] >= bmin) & (rec["r"] < bmax), axis=1)
            assert kept == len(ids) == inside.sum() and np.array_equal(ids, rec["id"][inside]) and max_id == rec["id"].max()
            parts.append(ids)
    assert len(parts[0]) and len(parts[1]) and np.array_equal(np.sort(np.concatenate(parts)), np.sort(rec["id"]))


@pytest.mark.gpu
def test_upload_records_errors():
    _, _, box = fcc_lattice(4)
    rng = np.random.default_rng(2)
    rec = synthetic_records(1000, "ICRV", 1, box, rng, outside=0.0)
    with ls1g.Container(box, 2.5) as c:
        with pytest.raises(ls1g.Ls1gError, match="ls1g_set_components first"):
            c.add_records(rec, "ICRV")
        c.set_components([ls1g.lj_component()])
        for wrong in (0, 2):                                                           # BinaryReader.cpp:222-228
            bad = rec.copy()
            bad["cid"][500] = wrong
            with pytest.raises(ls1g.Ls1gError, match="wrong component id"):
                c.add_records(bad, "ICRV")
            assert c.count() == (0, 0)
        assert c.add_records(rec, "ICRV")[0] == 1000
        assert c.add_records(rec[:0], "ICRV") == (0, 0) and c.count() == (0, 0) and c.download_records().shape == (0, 116)
    with ls1g.Container(box, 2.5, capacity=600) as c:
        c.set_components([ls1g.lj_component()])
        with pytest.raises(ls1g.Ls1gError, match="capacity"):
            c.add_records(rec, "ICRV")
