"""Phase-space formats: ASCII .inp, binary checkpoint, component XML.  CPU-only tests: round trips, the
reference's error behaviour, and - when the reference tree is mounted - its shipped example files against the golden
fixtures that the reference itself produced from them."""
import importlib
import os

import numpy as np
import pytest

from tests.helpers import load as load_golden

ls1g = importlib.import_module("ls1-mardyn_b200")
P = ls1g.phasespace
REF = "/root/reference/examples"


def synthetic(n=37, seed=5):
    rng = np.random.default_rng(seed)
    comps = [dict(lj=[[0, 0, 0, 0.012, 3.9e-5, 5.3, 0, 0], [0, 0, -2.4, 0.016, 3.1e-4, 5.6, 0, 0], [0, 0, 2.4, 0.016, 3.1e-4, 5.6, 0, 0]],
                  charges=[[0, 0, 0, 0, 21.2], [0, 0, -0.37, 0, -10.6], [0, 0, 0.37, 0, -10.6]], dipoles=[], quadrupoles=[], I=[0.18925, 0.18925, 0.0]),
             dict(lj=[[0, 1.4, 0, 0.016, 1.9e-4, 5.8, 12.5, 1]], charges=[], dipoles=[[0, 0, 0, 0, 1, 0, 0.96]],
                  quadrupoles=[[0.1, 0, 0, 0, 0, 1, 2.5]], I=[0.0, 0.0, 0.0])]
    q = rng.normal(size=(n, 4)); q /= np.linalg.norm(q, axis=1)[:, None]
    return P.PhaseSpace(time=1.25, box=np.array([30.0, 31.5, 29.0]), temperature=0.0012, components=comps, mixcoeff=[1.0, 0.95], eps_rf=1e10,
                        id=rng.permutation(n).astype(np.uint64) + 1, cid=rng.integers(0, 2, n).astype(np.uint32),
                        r=rng.uniform(0, 29, (n, 3)), v=rng.normal(size=(n, 3)), q=q, D=rng.normal(size=(n, 3)))


def same_molecules(a, b):
    for k in ("id", "cid", "r", "v", "q", "D"):
        assert np.array_equal(a[k], b[k]), k


def test_ascii_round_trip_is_bit_exact(tmp_path):
    ps = synthetic()
    P.write_ascii(tmp_path / "x.inp", ps)
    back = P.read_ascii(tmp_path / "x.inp")
    same_molecules(ps, back)
    assert back["time"] == ps["time"] and np.array_equal(back["box"], ps["box"]) and back["temperature"] == ps["temperature"]
    assert back["mixcoeff"] == ps["mixcoeff"] and back["eps_rf"] == ps["eps_rf"]
    assert back["components"] == [{k: [list(map(float, r)) for r in v] if k != "I" else list(map(float, v)) for k, v in c.items()} for c in ps["components"]]


def test_binary_round_trip_and_record_layout(tmp_path):
    ps = synthetic(n=101)
    P.write_binary(tmp_path / "cp.header.xml", tmp_path / "cp.dat", ps)
    assert os.path.getsize(tmp_path / "cp.dat") == 101 * 116          # FullMolecule::writeBinary: 8 + 4 + 13 * 8
    raw = open(tmp_path / "cp.dat", "rb").read()
    assert np.frombuffer(raw[:8], "<u8")[0] == ps["id"][0] and np.frombuffer(raw[8:12], "<u4")[0] == ps["cid"][0] + 1
    assert np.array_equal(np.frombuffer(raw[12:36], "<f8"), ps["r"][0])
    back = P.read_binary(tmp_path / "cp.header.xml", tmp_path / "cp.dat", n_components=2)
    same_molecules(ps, back)
    assert back["time"] == ps["time"] and np.allclose(back["box"], ps["box"], rtol=0, atol=0)


def test_empty_phase_space(tmp_path):
    ps = synthetic(n=0)
    P.write_ascii(tmp_path / "e.inp", ps)
    assert P.read_ascii(tmp_path / "e.inp").n == 0
    P.write_binary(tmp_path / "e.xml", tmp_path / "e.dat", ps)
    assert P.read_binary(tmp_path / "e.xml", tmp_path / "e.dat").n == 0


def test_reader_errors(tmp_path):
    ps = synthetic(n=5)
    P.write_ascii(tmp_path / "x.inp", ps)
    txt = open(tmp_path / "x.inp").read()
    open(tmp_path / "bad1.inp", "w").write(txt.replace("mardyn trunk", "moldy trunk"))
    with pytest.raises(ValueError):
        P.read_ascii(tmp_path / "bad1.inp")
    open(tmp_path / "bad2.inp", "w").write(txt.replace("NumberOfMolecules\t5", "NumberOfMolecules\t6"))
    with pytest.raises(ValueError):
        P.read_ascii(tmp_path / "bad2.inp")                                  # truncated molecule block
    bad = synthetic(n=5); bad["cid"][:] = 7
    P.write_ascii(tmp_path / "bad3.inp", bad)
    with pytest.raises(ValueError):
        P.read_ascii(tmp_path / "bad3.inp")                                  # component id out of range (ASCIIReader.cpp:386-393)
    P.write_binary(tmp_path / "h.xml", tmp_path / "d.dat", ps)
    open(tmp_path / "d.dat", "ab").truncate(116 * 4)
    with pytest.raises(ValueError):
        P.read_binary(tmp_path / "h.xml", tmp_path / "d.dat")
    open(tmp_path / "h2.xml", "w").write(open(tmp_path / "h.xml").read().replace("ICRVQD", "XYZ"))
    with pytest.raises(ValueError):
        P.read_binary(tmp_path / "h2.xml", tmp_path / "d.dat")


def test_component_xml_round_trip(tmp_path):
    ps = synthetic()
    open(tmp_path / "c.xml", "w").write(P.components_xml(ps["components"]))
    back = P.read_components_xml(tmp_path / "c.xml")
    for a, b in zip(ps["components"], back):
        for k in ("charges", "dipoles", "quadrupoles"):
            assert np.array_equal(np.array(a[k]).reshape(-1), np.array(b[k]).reshape(-1))
        la, lb = np.array(a["lj"]), np.array(b["lj"])
        assert np.array_equal(la[:, :6], lb[:, :6]) and np.array_equal(la[:, 7] != 0, lb[:, 7] != 0)
        assert list(a["I"]) == list(b["I"])


def test_abi_components_follow_component_cpp():
    ps = synthetic()
    abi = P.to_abi_components(ps["components"], lj_cutoff=10.0)
    c0, c1 = abi
    assert c0["mass"] == 0.012 + 0.016 + 0.016 and c0["rot_dof"] == 2                    # linear molecule: I33 = 0
    assert c0["inertia"][:2] == [0.18925, 0.18925] and c0["inertia"][2] == 0.0          # positive file values override
    assert c0["lj"][0, 6] == 0.0
    s6 = (5.8 / 12.5) ** 6                                                               # own tcutoff of the ASCII header wins
    assert c1["lj"][0, 6] == pytest.approx(24 * 1.9e-4 * (s6 - s6 * s6), rel=1e-15)
    assert c1["rot_dof"] == 2 and c1["inertia"][0] == pytest.approx(0.016 * 1.4 ** 2)    # one massive site on the y axis


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference tree not mounted")
@pytest.mark.parametrize("fixture,inp,comp", [("ex_CO2", "CO2/370K_15mol_l/CO2_370K_15mol_l.inp", "CO2/components.xml"),
                                               ("ex_EOX", "EOX/600K_15mol_l/EOX_600K_15mol_l.inp", "EOX/components.xml"),
                                               ("ex_Argon", "Argon/200K_18mol_l/Argon_200K_18mol_l.inp", None)])
def test_reference_examples_match_the_golden_dumps(fixture, inp, comp):
    """the golden fixtures hold what the REFERENCE read from these very files (ref_driver dump): same molecules, and the
    same component tables whether they come from the .inp header or from components.xml"""
    d = load_golden(fixture)
    ps = P.read_ascii(os.path.join(REF, inp))
    order, gorder = np.argsort(ps["id"]), np.argsort(d["in_id"])
    assert np.array_equal(ps["id"][order], d["in_id"][gorder].astype(np.uint64))
    assert np.array_equal(ps["cid"][order], d["in_cid"][gorder].astype(np.uint32))
    for k, g in (("r", "in_r"), ("v", "in_v"), ("D", "in_L")):
        assert np.array_equal(ps[k][order], d[g][gorder]), k
    q = ps["q"][order]                       # the reference's Molecule constructor normalises the quaternion it is given
    assert np.allclose(q / np.linalg.norm(q, axis=1)[:, None], d["in_q"][gorder], rtol=0, atol=4e-16)
    assert np.array_equal(ps["box"], d["box"])
    tables = [P.to_abi_components(ps["components"], lj_cutoff=float(d["rc"][1]))]
    if comp:
        tables.append(P.to_abi_components(P.read_components_xml(os.path.join(REF, comp)), lj_cutoff=float(d["rc"][1])))
    gold = ls1g.components_from_tables(d["comp_counts"], d["comp_scalars"], d["lj_sites"], d["charge_sites"], d["dipole_sites"], d["quadrupole_sites"])
    for abi in tables:
        assert len(abi) == len(gold)
        for a, g in zip(abi, gold):
            assert a["mass"] == pytest.approx(g["mass"], rel=1e-15) and a["rot_dof"] == g["rot_dof"]
            assert np.allclose(a["inertia"], g["inertia"], rtol=1e-15, atol=0)
            for k in ("lj", "charge", "dipole", "quadrupole"):
                assert np.allclose(a[k], g[k], rtol=1e-15, atol=0), k
