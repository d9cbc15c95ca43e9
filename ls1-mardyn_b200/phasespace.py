"""Phase-space and checkpoint formats of ls1-MarDyn on the device boundary.

Host-side mirror of the reference's readers/writers for the formats that feed the hot path:

  ASCII ``.inp``         io/ASCIIReader.cpp:76-262 (header: currentTime, Length, Temperature, NumberOfComponents block with
                         the site tables, mixing coefficients, epsilonRF) and :264-440 (NumberOfMolecules, MoleculeFormat
                         ICRVQD | ICRV | IRV records)
  binary checkpoint      io/BinaryReader.cpp + FullMolecule::writeBinary (FullMolecule.cpp:451-474): packed little-endian
                         records  id u64 | cid u32 | r 3xf64 | v 3xf64 | q 4xf64 | D 3xf64  = 116 bytes, next to an XML
                         header (io/CheckpointWriter.cpp, BinaryReader.cpp:44-100: number, format type="ICRVQD", time, length)
  component XML          molecules/Component.cpp readXML + Site.h (LJ126, Charge, Dipole, Quadrupole sites, moments of inertia)

``load`` / ``store`` move a whole phase space between a file and a device-resident Container in one upload / download
(pinned staging inside the C ABI): the molecule records never exist as per-molecule host objects.
Units and conventions are the reference's: component ids are 1-based in the files and 0-based on the device; D is the
angular momentum; the quaternion is (w, x, y, z).
"""
import os
import xml.etree.ElementTree as ET

import numpy as np

# FullMolecule::writeBinary record (FullMolecule.cpp:451-474)
BINARY_RECORD = np.dtype([("id", "<u8"), ("cid", "<u4"), ("r", "<f8", 3), ("v", "<f8", 3), ("q", "<f8", 4), ("D", "<f8", 3)])
assert BINARY_RECORD.itemsize == 116


class PhaseSpace(dict):
    """keys: time, box[3], temperature, components (list of dicts: lj, charges, dipoles, quadrupoles, I[3]), mixcoeff,
    eps_rf, id u64[n], cid u32[n] (0-based), r[n,3], v[n,3], q[n,4], D[n,3]"""

    @property
    def n(self):
        return len(self["id"])


# ----------------------------------------------------------------------------------------------------------
# ASCII (.inp)
# ----------------------------------------------------------------------------------------------------------
def _tokens(path):
    with open(path, "r") as f:
        for line in f:
            if line.lstrip().startswith("#"):         # comment lines ("# rho ...") are skipped by the reference's tokenizer too
                continue
            for t in line.split():
                yield t


def read_ascii(path):
    """ASCIIReader::readPhaseSpaceHeader + readPhaseSpace for a single-file phase space (header and molecules in one file)."""
    ps = PhaseSpace(time=0.0, temperature=None, components=[], mixcoeff=[], eps_rf=None)
    it = _tokens(path)
    tok = next(it)
    if tok != "mardyn":
        raise ValueError("%s: not a mardyn phase space file (first token %r)" % (path, tok))
    next(it)                      # "trunk"
    next(it)                      # version date (ASCIIReader.cpp:81-90 requires >= 20080701)
    fmt, n = "ICRVQD", None
    for tok in it:
        if tok in ("currentTime", "t"):
            ps["time"] = float(next(it))
        elif tok in ("Temperature", "T"):
            ps["temperature"] = float(next(it))
        elif tok in ("Length", "L"):
            ps["box"] = np.array([float(next(it)) for _ in range(3)])
        elif tok in ("NumberOfComponents", "C"):
            nc = int(next(it))
            for _ in range(nc):
                nlj, nch, ndi, nqu, nter = (int(next(it)) for _ in range(5))
                if nter:
                    raise ValueError("tersoff no longer supported")          # ASCIIReader.cpp:163-167
                rd = lambda k: [float(next(it)) for _ in range(k)]
                comp = dict(lj=[rd(8) for _ in range(nlj)],                 # x y z m eps sigma tcutoff do_shift
                            charges=[rd(5) for _ in range(nch)],            # x y z m q
                            dipoles=[rd(7) for _ in range(ndi)],            # x y z ex ey ez |mu|
                            quadrupoles=[rd(7) for _ in range(nqu)],        # x y z ex ey ez |Q|
                            I=rd(3))
                ps["components"].append(comp)
            for i in range(1, nc):
                for _ in range(i + 1, nc + 1):
                    ps["mixcoeff"] += [float(next(it)), float(next(it))]     # xi, eta
            ps["eps_rf"] = float(next(it))
        elif tok in ("NumberOfMolecules", "N"):
            n = int(next(it))
        elif tok in ("MoleculeFormat", "M"):
            fmt = next(it)
            break
        elif tok in ("ThermostatTemperature", "ThT", "h", "ComponentThermostat", "CT", "o"):
            next(it); next(it)
        elif tok in ("Undirected", "U"):
            next(it)
        elif tok in ("HeatCapacity", "cv", "I"):
            next(it); next(it); next(it)
        else:
            raise ValueError("%s: invalid token %r in the header" % (path, tok))
    if n is None:
        raise ValueError("%s: NumberOfMolecules missing" % path)
    width = {"ICRVQD": 15, "ICRVQDV": 18, "ICRV": 8, "IRV": 7}.get(fmt)
    if width is None:
        raise ValueError("unknown MoleculeFormat %r" % fmt)
    flat = np.array([float(t) for t in it], dtype=np.float64) if n else np.zeros(0)
    if flat.size != n * width:
        raise ValueError("%s: expected %d x %d molecule fields, found %d" % (path, n, width, flat.size))
    rec = flat.reshape(n, width)
    ps["id"] = rec[:, 0].astype(np.uint64)
    k = 1
    if fmt == "IRV":
        ps["cid"] = np.zeros(n, dtype=np.uint32)
    else:
        cid = rec[:, 1].astype(np.int64)
        if n and (cid.min() < 1 or cid.max() > max(len(ps["components"]), 1)):
            raise ValueError("%s: component id out of range" % path)          # ASCIIReader.cpp:386-393
        ps["cid"] = (cid - 1).astype(np.uint32)
        k = 2
    ps["r"] = np.ascontiguousarray(rec[:, k:k + 3]); ps["v"] = np.ascontiguousarray(rec[:, k + 3:k + 6])
    if fmt.startswith("ICRVQD"):
        ps["q"] = np.ascontiguousarray(rec[:, k + 6:k + 10]); ps["D"] = np.ascontiguousarray(rec[:, k + 10:k + 13])
    else:                                                                      # ASCIIReader.cpp:357-364 defaults
        ps["q"] = np.tile(np.array([1.0, 0.0, 0.0, 0.0]), (n, 1)); ps["D"] = np.zeros((n, 3))
    return ps


def write_ascii(path, ps):
    """The layout io/CheckpointWriter (ASCII branch) / Domain::writeCheckpointHeader produce, readable by read_ascii and by
    the reference's ASCIIReader."""
    with open(path, "w") as f:
        f.write("mardyn trunk 20120726\n currentTime\t%.17g\n Length\t%.17g %.17g %.17g\n" % ((ps.get("time", 0.0),) + tuple(ps["box"])))
        if ps.get("temperature") is not None:
            f.write(" Temperature\t%.17g\n" % ps["temperature"])
        comps = ps["components"]
        f.write(" NumberOfComponents\t%d\n" % len(comps))
        for c in comps:
            f.write("%d\t%d\t%d\t%d\t0\n" % (len(c["lj"]), len(c["charges"]), len(c["dipoles"]), len(c["quadrupoles"])))
            for group in ("lj", "charges", "dipoles", "quadrupoles"):
                for row in c[group]:
                    f.write(" ".join("%.17g" % x for x in row) + "\n")
            f.write("%.17g %.17g %.17g\n" % tuple(c["I"]))
        mc = list(ps.get("mixcoeff", []))
        if mc:
            f.write(" ".join("%.17g" % x for x in mc) + "\n")
        f.write("%.17g\n" % (ps["eps_rf"] if ps.get("eps_rf") is not None else 1e10))
        n = len(ps["id"])
        f.write(" NumberOfMolecules\t%d\n MoleculeFormat\tICRVQD\n" % n)
        out = np.empty((n, 15))
        out[:, 0] = ps["id"]; out[:, 1] = ps["cid"].astype(np.int64) + 1
        out[:, 2:5] = ps["r"]; out[:, 5:8] = ps["v"]; out[:, 8:12] = ps["q"]; out[:, 12:15] = ps["D"]
        np.savetxt(f, out, fmt=["%d", "%d"] + ["%.17g"] * 13, delimiter="\t")


# ----------------------------------------------------------------------------------------------------------
# binary checkpoint (header XML + packed records)
# ----------------------------------------------------------------------------------------------------------
_BIN_DTYPES = {
    "ICRVQD": BINARY_RECORD,
    "ICRV": np.dtype([("id", "<u8"), ("cid", "<u4"), ("r", "<f8", 3), ("v", "<f8", 3)]),
    "IRV": np.dtype([("id", "<u8"), ("r", "<f8", 3), ("v", "<f8", 3)]),
}


def read_binary(header_path, data_path, n_components=None):
    """BinaryReader::readPhaseSpaceHeader (BinaryReader.cpp:79-119) + readPhaseSpace (:121-260).  The component tables are
    not part of a binary checkpoint (the reference takes them from the XML configuration)."""
    root = ET.parse(header_path).getroot()
    if root.tag != "mardyn":
        raise ValueError("Not a valid MarDyn XML input file.")
    hi = root.find("headerinfo")
    try:
        time = float(hi.find("time").text)
        box = np.array([float(hi.find("length/" + a).text) for a in "xyz"])
        n = int(hi.find("number").text)
        fmt = hi.find("format").attrib["type"]
    except (AttributeError, KeyError, TypeError, ValueError):
        raise ValueError("Content of file: '%s' corrupted!" % header_path)
    if fmt not in _BIN_DTYPES:
        raise ValueError("Not a valid molecule format: %s" % fmt)
    rec = np.fromfile(data_path, dtype=_BIN_DTYPES[fmt], count=n)
    if len(rec) != n:
        raise ValueError("%s: %d records expected, %d found" % (data_path, n, len(rec)))
    ps = PhaseSpace(time=time, box=box, temperature=None, components=[], mixcoeff=[], eps_rf=None)
    ps["id"] = np.ascontiguousarray(rec["id"])
    if fmt == "IRV":
        ps["cid"] = np.zeros(n, dtype=np.uint32)                     # componentid = 1 (BinaryReader.cpp:207)
    else:
        cid = rec["cid"].astype(np.int64)
        if n and (cid.min() < 1 or (n_components is not None and cid.max() > n_components)):
            raise ValueError("Molecule has wrong componentid")        # BinaryReader.cpp:225-231
        ps["cid"] = (cid - 1).astype(np.uint32)
    ps["r"] = np.ascontiguousarray(rec["r"]); ps["v"] = np.ascontiguousarray(rec["v"])
    if fmt == "ICRVQD":
        ps["q"] = np.ascontiguousarray(rec["q"]); ps["D"] = np.ascontiguousarray(rec["D"])
    else:
        ps["q"] = np.tile(np.array([1.0, 0.0, 0.0, 0.0]), (n, 1)); ps["D"] = np.zeros((n, 3))
    return ps


def write_binary(header_path, data_path, ps):
    """Domain::writeCheckpointHeaderXML (Domain.cpp:572-595; FORMAT_SCI_MAX_DIGITS_WIDTH_21 = %21.15e, utils/FileUtils.h:17, so the box
    lengths keep 15 digits as in the reference) + FullMolecule::writeBinary per molecule.  Byte-identical to the reference's files
    (tests/golden/cp_binary_MultiComponent.*)."""
    n = len(ps["id"])
    with open(header_path, "w") as f:
        f.write("<?xml version='1.0' encoding='UTF-8'?>\n<mardyn version=\"20100525\" >\n\t<headerinfo>\n")
        f.write("\t\t<time>%21.15e</time>\n\t\t<length>\n\t\t\t<x>%21.15e</x> <y>%21.15e</y> <z>%21.15e</z>\n\t\t</length>\n"
                % ((ps.get("time", 0.0),) + tuple(ps["box"])))
        f.write("\t\t<number>%d</number>\n\t\t<format type=\"ICRVQD\"/>\n\t</headerinfo>\n</mardyn>\n" % n)
    rec = np.empty(n, dtype=BINARY_RECORD)
    rec["id"] = ps["id"]; rec["cid"] = ps["cid"].astype(np.uint32) + 1
    rec["r"] = ps["r"]; rec["v"] = ps["v"]; rec["q"] = ps["q"]; rec["D"] = ps["D"]
    rec.tofile(data_path)


# ----------------------------------------------------------------------------------------------------------
# components: ASCII header tables <-> XML <-> the C ABI's ls1g_component rows
# ----------------------------------------------------------------------------------------------------------
def components_xml(components, names=None):
    """<components><moleculetype> text as Component::readXML reads it (Component.cpp:40-102, Site.h readXML)."""
    out = ["<?xml version='1.0' encoding='UTF-8'?>", "<components version=\"20100525\" >"]
    for ci, c in enumerate(components):
        out.append("\t<moleculetype id=\"%d\" name=\"%s\">" % (ci + 1, (names or {}).get(ci, "comp%d" % (ci + 1))))
        sid = 0
        for x, y, z, m, eps, sig, _tc, shift in c["lj"]:
            sid += 1
            out.append("\t\t<site type=\"LJ126\" id=\"%d\" ><coords> <x>%.17g</x> <y>%.17g</y> <z>%.17g</z> </coords><mass>%.17g</mass>"
                       "<sigma>%.17g</sigma><epsilon>%.17g</epsilon><shifted>%d</shifted></site>" % (sid, x, y, z, m, sig, eps, int(shift != 0)))
        for x, y, z, m, q in c["charges"]:
            sid += 1
            out.append("\t\t<site type=\"Charge\" id=\"%d\" ><coords> <x>%.17g</x> <y>%.17g</y> <z>%.17g</z> </coords><mass>%.17g</mass>"
                       "<charge>%.17g</charge></site>" % (sid, x, y, z, m, q))
        for x, y, z, ex, ey, ez, mu in c["dipoles"]:
            sid += 1
            out.append("\t\t<site type=\"Dipole\" id=\"%d\" ><coords> <x>%.17g</x> <y>%.17g</y> <z>%.17g</z> </coords>"
                       "<dipolemoment><x>%.17g</x> <y>%.17g</y> <z>%.17g</z><abs>%.17g</abs></dipolemoment></site>" % (sid, x, y, z, ex, ey, ez, mu))
        for x, y, z, ex, ey, ez, Q in c["quadrupoles"]:
            sid += 1
            out.append("\t\t<site type=\"Quadrupole\" id=\"%d\" ><coords> <x>%.17g</x> <y>%.17g</y> <z>%.17g</z> </coords>"
                       "<quadrupolemoment><x>%.17g</x> <y>%.17g</y> <z>%.17g</z><abs>%.17g</abs></quadrupolemoment></site>" % (sid, x, y, z, ex, ey, ez, Q))
        out.append("\t\t<momentsofinertia rotaxes=\"xyz\" ><Ixx>%.17g</Ixx><Iyy>%.17g</Iyy><Izz>%.17g</Izz></momentsofinertia>" % tuple(c["I"]))
        out.append("\t</moleculetype>")
    out.append("</components>")
    return "\n".join(out) + "\n"


def read_components_xml(path):
    """the inverse of components_xml, for the reference's own components.xml files"""
    comps = []
    f = lambda e, tag: float(e.find(tag).text)
    for mt in ET.parse(path).getroot().iter("moleculetype"):
        c = dict(lj=[], charges=[], dipoles=[], quadrupoles=[], I=[0.0, 0.0, 0.0])
        for s in mt.findall("site"):
            xyz = [f(s, "coords/" + a) for a in "xyz"]
            t = s.attrib["type"]
            if t == "LJ126":
                sh = s.find("shifted")
                shifted = sh is not None and sh.text.strip().lower() in ("1", "true", "yes")
                c["lj"].append(xyz + [f(s, "mass"), f(s, "epsilon"), f(s, "sigma"), 0.0, 1.0 if shifted else 0.0])
            elif t == "Charge":
                c["charges"].append(xyz + [f(s, "mass"), f(s, "charge")])
            elif t == "Dipole":
                c["dipoles"].append(xyz + [f(s, "dipolemoment/" + a) for a in "xyz"] + [f(s, "dipolemoment/abs")])
            elif t == "Quadrupole":
                c["quadrupoles"].append(xyz + [f(s, "quadrupolemoment/" + a) for a in "xyz"] + [f(s, "quadrupolemoment/abs")])
            else:
                raise ValueError("unknown site type %r" % t)
        moi = mt.find("momentsofinertia")
        if moi is not None:
            c["I"] = [f(moi, "Ixx"), f(moi, "Iyy"), f(moi, "Izz")]
        comps.append(c)
    return comps


def to_abi_components(components, lj_cutoff=None):
    """ASCII-header / XML component tables -> the dicts Container.set_components takes (ls1g_component rows).
    Mass, principal moments and rotational degrees of freedom follow Component::updateMassInertia (Component.cpp:125-166):
    m = sum of the LJ and charge site masses, I_d from the site positions, rot_dof = 3 - #(I_d == 0); a positive I in the
    file overrides the computed moment (ASCIIReader.cpp:205-208, Component.cpp:93-95).  shift6 as Component::addLJcenter
    (Component.cpp:105-117): the ASCII header carries its own cutoff per site (tcutoff), the XML form uses rc_LJ."""
    out = []
    for c in components:
        m, I = 0.0, [0.0, 0.0, 0.0]
        for row in list(c["lj"]) + list(c["charges"]):
            x, y, z, ms = row[0], row[1], row[2], row[3]
            m += ms
            I[0] += ms * (y * y + z * z); I[1] += ms * (x * x + z * z); I[2] += ms * (x * x + y * y)
        rot_dof = 3 - sum(1 for d in range(3) if I[d] == 0.0)
        for d in range(3):
            if c["I"][d] > 0.0:
                I[d] = float(c["I"][d])
        lj = []
        for x, y, z, ms, eps, sig, tcut, do_shift in c["lj"]:
            shift6 = 0.0
            if do_shift != 0:
                rc = tcut if tcut > 0 else lj_cutoff
                if not rc:
                    raise ValueError("shifted LJ site without a cutoff")
                s2 = sig * sig / (rc * rc); s6 = s2 * s2 * s2
                shift6 = 24.0 * eps * (s6 - s6 * s6)
            lj.append([x, y, z, ms, eps, sig, shift6])
        out.append(dict(mass=m, inertia=I, rot_dof=rot_dof, lj=np.array(lj, dtype=np.float64).reshape(-1, 7),
                        charge=np.array(c["charges"], dtype=np.float64).reshape(-1, 5),
                        dipole=np.array(c["dipoles"], dtype=np.float64).reshape(-1, 7),
                        quadrupole=np.array(c["quadrupoles"], dtype=np.float64).reshape(-1, 7)))
    return out


# ----------------------------------------------------------------------------------------------------------
# file <-> device
# ----------------------------------------------------------------------------------------------------------
def load(container, ps, lj_cutoff=None, set_components=True):
    """Upload a phase space into a Container in one call (InputBase::readPhaseSpace -> ParticleContainer::addParticle for
    every molecule in the reference, ASCIIReader.cpp:395-430).  Molecules outside this rank's box are dropped by the
    container's own exchange, as the reference's readers do per rank."""
    if set_components and ps["components"]:
        container.set_components(to_abi_components(ps["components"], lj_cutoff), np.asarray(ps.get("mixcoeff", []), dtype=np.float64))
    container.add_particles(ps["id"], ps["cid"], ps["r"], ps["v"], q=ps["q"], L=ps["D"])
    return ps.n


def store(container, template=None, time=0.0):
    """Download the owned molecules of a Container as a PhaseSpace (CheckpointWriter::endStep -> Domain::writeCheckpoint,
    sorted by molecule id so that the file does not depend on the cell order)."""
    out = container.download(("id", "cid", "r", "v", "q", "L"))
    order = np.argsort(out["id"], kind="stable")
    ps = PhaseSpace(time=time, temperature=None, components=[], mixcoeff=[], eps_rf=None)
    if template is not None:
        for k in ("box", "temperature", "components", "mixcoeff", "eps_rf"):
            ps[k] = template.get(k)
    ps["id"] = out["id"][order]; ps["cid"] = out["cid"][order].astype(np.uint32)
    ps["r"] = out["r"][order]; ps["v"] = out["v"][order]; ps["q"] = out["q"][order]; ps["D"] = out["L"][order]
    return ps
