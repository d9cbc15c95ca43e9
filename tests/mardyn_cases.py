"""Self-contained ls1-MarDyn run directories (XML config + ASCII phase space) built from the golden fixtures, so that the
reference binary (oracle/_ref/MarDyn) and the same binary with the B200 container (ls1-mardyn_b200/host/_build/MarDynGpu)
can be run side by side on a box that has no /root/reference.  Test infrastructure."""
import importlib
import os
import re
import subprocess

import numpy as np

from tests.helpers import load

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_EXE = os.path.join(ROOT, "oracle", "_ref", "MarDyn")
GPU_EXE = os.path.join(ROOT, "ls1-mardyn_b200", "host", "_build", "MarDynGpu")
# <temperature> of the shipped example configs (examples/*/config.xml); the fixtures hold everything else
TEMPERATURE = {"ex_Argon": 0.000633363365, "ex_CO2": 0.00117172223, "ex_EOX": 0.0019000901}

CONFIG = """<?xml version='1.0' encoding='UTF-8'?>
<mardyn version="20100525">
  <refunits type="SI"><length unit="nm">0.0529177</length><mass unit="u">1000</mass><energy unit="eV">27.2126</energy></refunits>
  <simulation type="MD">
    <integrator type="{integrator}"><timestep unit="reduced">{dt!r}</timestep></integrator>
    <run><currenttime>0</currenttime>{equilibration}<production><steps>{steps}</steps></production></run>
    <ensemble type="NVT">
      <temperature unit="reduced">{temperature!r}</temperature>
      <domain type="box"><lx>{lx!r}</lx><ly>{ly!r}</ly><lz>{lz!r}</lz></domain>
      <components>
{moleculetypes}
      </components>
      <phasespacepoint><file type="ASCII">{inp}</file></phasespacepoint>
    </ensemble>
    <algorithm>
      <parallelisation type="DomainDecomposition"></parallelisation>
      <datastructure type="{container}"><cellsInCutoffRadius>1</cellsInCutoffRadius>{container_options}</datastructure>
      <cutoffs type="CenterOfMass"><defaultCutoff unit="reduced">{rc!r}</defaultCutoff><radiusLJ unit="reduced">{rclj!r}</radiusLJ></cutoffs>
      <electrostatic type="ReactionField"><epsilon>{eps_rf!r}</epsilon></electrostatic>
    </algorithm>
    <output>
      <outputplugin name="ResultWriter"><writefrequency>1</writefrequency><writeprecision>15</writeprecision><outputprefix>{prefix}</outputprefix></outputplugin>{output_plugins}
    </output>
  </simulation>
</mardyn>
"""


def phase_space_of(d):
    """golden fixture -> ls1g.phasespace.PhaseSpace (component tables split per component by comp_counts)"""
    P = importlib.import_module("ls1-mardyn_b200").phasespace
    comps, o = [], [0, 0, 0, 0]
    for k, cnt in enumerate(d["comp_counts"]):
        nlj, nch, ndi, nqu = (int(x) for x in cnt[:4])
        lj = [[x, y, z, m, eps, sig, 0.0, float(shift != 0.0)] for x, y, z, m, eps, sig, shift in d["lj_sites"][o[0]:o[0] + nlj]]
        comps.append(dict(lj=lj, charges=d["charge_sites"][o[1]:o[1] + nch].tolist(), dipoles=d["dipole_sites"][o[2]:o[2] + ndi].tolist(),
                          quadrupoles=d["quadrupole_sites"][o[3]:o[3] + nqu].tolist(), I=d["comp_scalars"][k][1:4].tolist()))
        o = [o[0] + nlj, o[1] + nch, o[2] + ndi, o[3] + nqu]
    return P.PhaseSpace(time=0.0, box=np.asarray(d["box"], dtype=np.float64), temperature=None, components=comps,
                        mixcoeff=list(d["mixcoeff"]), eps_rf=float(d["epsRF"][0]),
                        id=d["in_id"].astype(np.uint64), cid=d["in_cid"].astype(np.uint32), r=d["in_r"], v=d["in_v"], q=d["in_q"], D=d["in_L"])


def write_case(dirname, name, steps, variant, temperature=None, replicate=1, container_options="", output_plugins="", equilibration=None):
    """variant: 'cpu' (LinkedCells + Leapfrog), 'fused' / 'operators' (GpuLinkedCells + GpuLeapfrog); output_plugins: extra
    <outputplugin> elements; equilibration: <run><equilibration><steps> (sampling plugins start after it, default 20000).  Returns the config path."""
    P = importlib.import_module("ls1-mardyn_b200").phasespace
    d = load(name)
    ps = phase_space_of(d)
    if replicate > 1:
        ps = replicate_box(ps, replicate)
    inp = os.path.join(dirname, name + ".inp")
    if not os.path.exists(inp):
        P.write_ascii(inp, ps)
    xml = P.components_xml(ps["components"])
    moleculetypes = "\n".join(l for l in xml.splitlines() if not l.startswith("<?xml") and "components" not in l)
    gpu = variant != "cpu"                    # variant may carry a suffix after "-" to keep file names apart (e.g. "fused-mirror")
    cfg = os.path.join(dirname, "%s_%s.xml" % (name, variant))
    with open(cfg, "w") as f:
        f.write(CONFIG.format(integrator="GpuLeapfrog" if gpu else "Leapfrog", container="GpuLinkedCells" if gpu else "LinkedCells",
                              container_options=("<mode>%s</mode>" % variant.split("-")[0] + container_options) if gpu else "", dt=float(d["dt"][0]), steps=steps,
                              temperature=float(TEMPERATURE[name] if temperature is None else temperature),
                              lx=float(ps["box"][0]), ly=float(ps["box"][1]), lz=float(ps["box"][2]), moleculetypes=moleculetypes,
                              inp=os.path.basename(inp), rc=float(d["rc"][0]), rclj=float(d["rc"][1]), eps_rf=float(d["epsRF"][0]),
                              prefix="%s_%s" % (name, variant), output_plugins=output_plugins,
                              equilibration="" if equilibration is None else "<equilibration><steps>%d</steps></equilibration>" % equilibration))
    return cfg


def write_fcc_case(dirname, cells, steps, gpu, tag):
    """bench.py's workload (1CLJ fcc from the reference's own MultiObjectGenerator) with a ResultWriter; returns the config path"""
    import bench
    xml = bench.mardyn_xml(steps, cells, gpu)
    xml = xml.replace("<output></output>", "<output><outputplugin name=\"ResultWriter\"><writefrequency>1</writefrequency><writeprecision>15</writeprecision>"
                                           "<outputprefix>%s</outputprefix></outputplugin></output>" % tag)
    cfg = os.path.join(dirname, tag + ".xml")
    with open(cfg, "w") as f:
        f.write(xml)
    return cfg


def replicate_box(ps, k):
    """ReplicaGenerator-style k^3 replication of a periodic box (the examples replicated for throughput runs)"""
    P = importlib.import_module("ls1-mardyn_b200").phasespace
    n = len(ps["id"])
    shifts = np.array([[i, j, l] for i in range(k) for j in range(k) for l in range(k)], dtype=np.float64) * ps["box"]
    rep = lambda a: np.tile(a, (k ** 3,) + (1,) * (a.ndim - 1))
    out = P.PhaseSpace(ps)
    out["box"] = ps["box"] * k
    out["r"] = (ps["r"][None, :, :] + shifts[:, None, :]).reshape(-1, 3)
    out["id"] = (np.arange(k ** 3, dtype=np.uint64)[:, None] * np.uint64(n) + ps["id"][None, :]).reshape(-1)
    for key in ("cid", "v", "q", "D"):
        out[key] = rep(ps[key])
    return out


def read_checkpoint_molecules(path):
    """molecule records (ICRVQD) of a checkpoint written by Domain::writeCheckpoint; the component header the reference
    writes there (Component::write) is not the layout its ASCIIReader reads, so only the records are parsed.  Sorted by id."""
    with open(path) as f:
        lines = f.readlines()
    k = next(i for i, l in enumerate(lines) if l.split()[:1] == ["MoleculeFormat"])
    if lines[k].split()[1] != "ICRVQD":
        raise ValueError("unexpected molecule format in " + path)
    a = np.loadtxt(lines[k + 1:], ndmin=2)
    a = a[np.argsort(a[:, 0])]
    return dict(id=a[:, 0].astype(np.uint64), cid=a[:, 1].astype(np.uint32) - 1, r=a[:, 2:5], v=a[:, 5:8], q=a[:, 8:12], D=a[:, 12:15])


def run(exe, cfg, steps, threads=None, timeout=1800, check=True):
    """runs MarDyn / MarDynGpu in the config's directory; returns dict(log, res[steps+1, 9], loop_s, final (PhaseSpace));
    check=False: a failed run returns dict(log, returncode) instead of raising"""
    P = importlib.import_module("ls1-mardyn_b200").phasespace
    cwd = os.path.dirname(cfg)
    env = dict(os.environ)
    if threads:
        env["OMP_NUM_THREADS"] = str(threads)
    p = subprocess.run([exe, os.path.basename(cfg), "--steps", str(steps), "--final-checkpoint=1"], cwd=cwd, env=env,
                       capture_output=True, text=True, timeout=timeout)
    log = p.stdout + p.stderr
    if (p.returncode != 0 or "Exit code would have been" in log) and not check:
        return dict(log=log, returncode=p.returncode)
    if p.returncode != 0 or "Exit code would have been" in log:
        raise RuntimeError("%s %s failed (%d):\n%s" % (os.path.basename(exe), cfg, p.returncode, log[-3000:]))
    prefix = os.path.splitext(os.path.basename(cfg))[0]
    res = np.loadtxt(os.path.join(cwd, prefix + ".res"), comments="#", ndmin=2)
    m = re.search(r"Computation in main loop took:\s*([0-9.eE+-]+)", log)
    out = dict(log=log, res=res, loop_s=float(m.group(1)) if m else None)
    cp = os.path.join(cwd, prefix + ".restart.dat")
    if os.path.exists(cp):
        out["final"] = read_checkpoint_molecules(cp)
    g = re.search(r"GpuLinkedCells: (\d+) device steps took ([0-9.eE+-]+) s; upload ([0-9.eE+-]+) s \((\d+) B\); (\d+) host mirror download\(s\) took ([0-9.eE+-]+) s \((\d+) B\); context creation ([0-9.eE+-]+) s", log)
    if g:
        out["gpu"] = dict(steps=int(g.group(1)), device_s=float(g.group(2)), upload_s=float(g.group(3)), upload_bytes=int(g.group(4)),
                          mirrors=int(g.group(5)), mirror_s=float(g.group(6)), mirror_bytes=int(g.group(7)), context_s=float(g.group(8)))
    return out
