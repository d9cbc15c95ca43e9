"""The drop-in itself: ls1-MarDyn's own executable and UNMODIFIED time loop (Simulation::simulate, src/Simulation.cpp:920-1214)
with <datastructure type="GpuLinkedCells"> + <integrator type="GpuLeapfrog"> (ls1-mardyn_b200/host/) against the same
executable with LinkedCells + Leapfrog, on the reference's shipped examples (initial states from the golden fixtures).
Compared per time step: U_pot, p, beta_trans, beta_rot as the reference's ResultWriter prints them (15 digits), and the
final checkpoint (r, v, q, D per molecule)."""
import os

import numpy as np
import pytest

from tests import mardyn_cases as M

STEPS = 20
needs_binaries = pytest.mark.skipif(not (os.path.exists(M.REF_EXE) and os.path.exists(M.GPU_EXE)),
                                    reason="oracle/_ref/MarDyn or host/_build/MarDynGpu not built (needs the reference tree at build time)")


def rel(a, b):
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))


@needs_binaries
def test_patched_binary_keeps_the_reference_path_bit_identical(tmp_path):
    """MarDynGpu = reference objects + two dispatch branches: with LinkedCells/Leapfrog it must reproduce MarDyn exactly"""
    a = M.run(M.REF_EXE, M.write_case(str(tmp_path), "ex_EOX", 5, "cpu"), 5, threads=1)
    os.rename(tmp_path / "ex_EOX_cpu.res", tmp_path / "first.res")
    b = M.run(M.GPU_EXE, M.write_case(str(tmp_path), "ex_EOX", 5, "cpu"), 5, threads=1)
    assert np.array_equal(a["res"], b["res"])
    for k in ("id", "r", "v", "q", "D"):
        assert np.array_equal(a["final"][k], b["final"][k]), k


@needs_binaries
def test_gpu_container_fails_loudly_without_a_device(tmp_path):
    """no CPU fallback: without a CUDA device the GPU container ends the run through the reference's fatal path"""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(RuntimeError, match="no CUDA device"):
        M.run(M.GPU_EXE, M.write_case(str(tmp_path), "ex_Argon", 2, "fused"), 2, threads=2)


WRITERS = ('<outputplugin name="CheckpointWriter"><type>ASCII</type><writefrequency>%d</writefrequency><outputprefix>%s</outputprefix>'
           '<incremental>1</incremental></outputplugin>'
           '<outputplugin name="XyzWriter"><writefrequency>%d</writefrequency><outputprefix>%s</outputprefix></outputplugin>'
           '<outputplugin name="SysMonOutput"><writefrequency>1</writefrequency><expression>procloadavg:loadavg1</expression></outputplugin>')


@needs_binaries
def test_mirror_frequency_follows_the_writers(tmp_path):
    """<mirrorfrequency> defaults to auto: gcd of the write frequencies of the plugins that read molecules (CheckpointWriter 6,
    XyzWriter 4 -> 2); ResultWriter and SysMonOutput only use Domain values and do not count.  Runs up to the first force pass on CPU."""
    import torch
    cfg = M.write_case(str(tmp_path), "ex_Argon", 2, "fused", output_plugins=WRITERS % (6, "cp", 4, "xyz"))
    out = M.run(M.GPU_EXE, cfg, 2, threads=2, check=False)
    assert "GpuLinkedCells: plugin CheckpointWriter reads molecules every 6 step(s)" in out["log"]
    assert "GpuLinkedCells: plugin XyzWriter reads molecules every 4 step(s)" in out["log"]
    assert "mirror frequency 2 (from the plugins' write frequencies)" in out["log"]
    assert "plugin ResultWriter" not in out["log"] and "plugin SysMonOutput" not in out["log"]
    if not torch.cuda.is_available():
        assert "no CUDA device" in out["log"]
    # an explicit number wins
    cfg = M.write_case(str(tmp_path), "ex_Argon", 2, "fused-explicit", container_options="<mirrorfrequency>0</mirrorfrequency>",
                       output_plugins=WRITERS % (6, "cp", 4, "xyz"))
    out = M.run(M.GPU_EXE, cfg, 2, threads=2, check=False)
    assert "mirror frequency 0, device" in out["log"]


@pytest.mark.gpu
@needs_binaries
def test_checkpoints_written_during_the_run_match_the_reference(tmp_path):
    """CheckpointWriter every 5 steps of a 10-step EOX run (NVT; the limiter of Domain.cpp:254-311 acts in the first steps): the
    writer reads the host cells, which the container refills on exactly those steps - both checkpoints equal the CPU reference's"""
    cpu_dir, gpu_dir = tmp_path / "cpu", tmp_path / "gpu"
    cpu_dir.mkdir(), gpu_dir.mkdir()
    M.run(M.REF_EXE, M.write_case(str(cpu_dir), "ex_EOX", 10, "cpu", output_plugins=WRITERS % (5, "cp", 5, "xyz")), 10)
    gpu = M.run(M.GPU_EXE, M.write_case(str(gpu_dir), "ex_EOX", 10, "fused", output_plugins=WRITERS % (5, "cp", 5, "xyz")), 10)
    assert "mirror frequency 5 (from the plugins' write frequencies)" in gpu["log"]
    assert gpu["gpu"]["mirrors"] == 2
    for k in (0, 1, 2):            # 0 = the start configuration (Simulation::prepare_start), written from the host phase
        a = M.read_checkpoint_molecules(str(cpu_dir / ("cp-%d.restart.dat" % k)))
        b = M.read_checkpoint_molecules(str(gpu_dir / ("cp-%d.restart.dat" % k)))
        ia, ib = np.argsort(a["id"]), np.argsort(b["id"])
        assert np.array_equal(a["id"][ia], b["id"][ib]) and len(ia) > 0
        for key in ("r", "v", "q", "D"):
            assert np.allclose(a[key][ia], b[key][ib], rtol=1e-7, atol=1e-9), (k, key)


BIN_REF = ('<outputplugin name="CheckpointWriter"><type>binary</type><writefrequency>%d</writefrequency><outputprefix>cp</outputprefix>'
           '</outputplugin>')
BIN_GPU = '<outputplugin name="GpuCheckpointWriter"><writefrequency>%d</writefrequency><outputprefix>cp</outputprefix></outputplugin>'
RECORD = np.dtype([("id", "<u8"), ("cid", "<u4"), ("r", "<f8", 3), ("v", "<f8", 3), ("q", "<f8", 4), ("D", "<f8", 3)])


def records_by_id(path):
    rec = np.fromfile(path, dtype=RECORD)
    return rec[np.argsort(rec["id"])]


@needs_binaries
def test_device_checkpoint_plugin_equals_the_reference_writer_on_a_host_container(tmp_path):
    """GpuCheckpointWriter with the reference's own LinkedCells (no device): same file names, same bytes as <CheckpointWriter
    type=binary> of the unmodified reference - the plugin's registration, XML and naming, on CPU"""
    cpu_dir, gpu_dir = tmp_path / "ref", tmp_path / "plugin"
    cpu_dir.mkdir(), gpu_dir.mkdir()
    M.run(M.REF_EXE, M.write_case(str(cpu_dir), "ex_CO2", 4, "cpu", output_plugins=BIN_REF % 2), 4, threads=1)
    M.run(M.GPU_EXE, M.write_case(str(gpu_dir), "ex_CO2", 4, "cpu", output_plugins=BIN_GPU % 2), 4, threads=1)
    names = sorted(f for f in os.listdir(cpu_dir) if f.startswith("cp-"))
    assert names == ["cp-%d.restart.%s" % (k, e) for k in (0, 1, 2) for e in ("dat", "header.xml")]
    for f in names:
        assert open(cpu_dir / f, "rb").read() == open(gpu_dir / f, "rb").read(), f


ASCII_REF = ('<outputplugin name="CheckpointWriter"><type>ASCII</type><writefrequency>%d</writefrequency><outputprefix>cp</outputprefix>'
             '</outputplugin>')
ASCII_GPU = ('<outputplugin name="GpuCheckpointWriter"><type>ASCII</type><writefrequency>%d</writefrequency><outputprefix>cp</outputprefix>'
             '</outputplugin>')


def ascii_header(path):
    text = open(path, "rb").read()
    return text[:text.index(b"\n", text.index(b"MoleculeFormat")) + 1]


@needs_binaries
def test_device_checkpoint_plugin_ascii_on_a_host_container(tmp_path):
    """<type>ASCII</type> with the reference's own LinkedCells: the bytes of <CheckpointWriter type=ASCII>"""
    cpu_dir, gpu_dir = tmp_path / "ref", tmp_path / "plugin"
    cpu_dir.mkdir(), gpu_dir.mkdir()
    M.run(M.REF_EXE, M.write_case(str(cpu_dir), "ex_EOX", 2, "cpu", output_plugins=ASCII_REF % 2), 2, threads=1)
    M.run(M.GPU_EXE, M.write_case(str(gpu_dir), "ex_EOX", 2, "cpu", output_plugins=ASCII_GPU % 2), 2, threads=1)
    names = sorted(f for f in os.listdir(cpu_dir) if f.startswith("cp-"))
    assert names == ["cp-0.restart.dat", "cp-1.restart.dat"]
    for f in names:
        assert open(cpu_dir / f, "rb").read() == open(gpu_dir / f, "rb").read(), f


@pytest.mark.gpu
@needs_binaries
@pytest.mark.parametrize("name,devices", [("ex_CO2", 1), ("ex_Argon", 2)])
def test_ascii_checkpoints_taken_from_the_device_records(name, devices, tmp_path):
    """GpuCheckpointWriter <type>ASCII</type>: the reference's header, then lines printed by the host threads from the device's packed
    records (no mirror step).  Header equal to the CPU reference's byte for byte; molecules against the CPU reference (file-wide
    tolerance) and against the same device trajectory written by the reference's ASCII writer on mirror steps (1e-12)"""
    opts = "<devices>%d</devices>" % devices if devices > 1 else ""
    dirs = {k: tmp_path / k for k in ("cpu", "records", "mirror")}
    for d in dirs.values():
        d.mkdir()
    M.run(M.REF_EXE, M.write_case(str(dirs["cpu"]), name, 7, "cpu", output_plugins=ASCII_REF % 3), 7)
    rec = M.run(M.GPU_EXE, M.write_case(str(dirs["records"]), name, 7, "fused", container_options=opts, output_plugins=ASCII_GPU % 3), 7)
    M.run(M.GPU_EXE, M.write_case(str(dirs["mirror"]), name, 7, "fused", container_options=opts, output_plugins=ASCII_REF % 3), 7)
    assert rec["gpu"]["mirrors"] == 1 and "2 record checkpoint(s) from the device" in rec["log"]
    for k in (0, 1, 2):
        f = "cp-%d.restart.dat" % k
        assert ascii_header(str(dirs["cpu"] / f)) == ascii_header(str(dirs["records"] / f))
        a, b, c = (M.read_checkpoint_molecules(str(dirs[w] / f)) for w in ("cpu", "records", "mirror"))
        assert len(a["id"]) == len(b["id"]) == len(c["id"]) > 0
        assert np.array_equal(a["id"], b["id"]) and np.array_equal(a["cid"], b["cid"]) and np.array_equal(b["id"], c["id"])
        for key in ("r", "v", "q", "D"):
            assert np.allclose(a[key], b[key], rtol=1e-7, atol=1e-9), (k, key)
            scale = max(np.abs(c[key]).max(), 1e-300)
            assert np.abs(b[key] - c[key]).max() <= 1e-12 * scale, (k, key)


@pytest.mark.gpu
@needs_binaries
@pytest.mark.parametrize("name,variant,devices,steps", [("ex_Argon", "fused", 1, 7), ("ex_CO2", "fused", 1, 7), ("ex_EOX", "fused", 1, 7),
                                                        ("ex_EOX", "operators", 1, 7), ("ex_CO2", "fused", 2, 7), ("ex_CO2", "fused", 1, 6)])
def test_checkpoints_taken_from_the_device_records(name, variant, devices, steps, tmp_path):
    """GpuCheckpointWriter every 3 steps of a 7-step NVT run (6 steps: the second checkpoint falls on the last step, whose state is
    already mirrored into the host cells, and goes through the reference's writer): the packed records come straight from the device (no mirror step, no
    host Molecule), with the step's velocity scaling - and for EOX the kinetic-energy limiter of Domain.cpp:254-311 - applied as the
    reference has by the time its writer runs.  Held to (a) the CPU reference's binary CheckpointWriter files and (b) the SAME device
    trajectory written through the reference's writer on mirror steps, which must give the same records to rounding."""
    opts = "<devices>%d</devices>" % devices if devices > 1 else ""
    dirs = {k: tmp_path / k for k in ("cpu", "records", "mirror")}
    for d in dirs.values():
        d.mkdir()
    M.run(M.REF_EXE, M.write_case(str(dirs["cpu"]), name, steps, "cpu", output_plugins=BIN_REF % 3), steps)
    rec = M.run(M.GPU_EXE, M.write_case(str(dirs["records"]), name, steps, variant, container_options=opts, output_plugins=BIN_GPU % 3), steps)
    mir = M.run(M.GPU_EXE, M.write_case(str(dirs["mirror"]), name, steps, variant, container_options=opts, output_plugins=BIN_REF % 3), steps)
    assert "mirror frequency 0 (from the plugins' write frequencies)" in rec["log"] and rec["gpu"]["mirrors"] == 1      # the final one only
    from_device = sum(1 for k in (3, 6) if k < steps)
    assert "%d record checkpoint(s) from the device" % from_device in rec["log"] and mir["gpu"]["mirrors"] == len({3, 6, steps})
    if name == "ex_EOX":
        assert "selective thermostat applied 0 time(s)" not in rec["log"]
        sel = lambda log: log.split("selective thermostat applied ")[1].split(" ")[0]
        assert sel(rec["log"]) == sel(mir["log"])                                                 # once per step, whoever asks first
    assert np.allclose(rec["res"], mir["res"], rtol=1e-12, atol=0)                               # taking checkpoints does not touch the trajectory
    for k in (0, 1, 2):            # 0 = the start configuration, written from the host phase through the reference's writer
        hdr = "cp-%d.restart.header.xml" % k
        assert open(dirs["cpu"] / hdr, "rb").read() == open(dirs["records"] / hdr, "rb").read()
        a, b, c = (records_by_id(str(dirs[w] / ("cp-%d.restart.dat" % k))) for w in ("cpu", "records", "mirror"))
        assert len(a) == len(b) == len(c) > 0
        assert np.array_equal(a["id"], b["id"]) and np.array_equal(a["cid"], b["cid"]) and np.array_equal(b["id"], c["id"])
        for key in ("r", "v", "q", "D"):
            assert np.allclose(a[key], b[key], rtol=1e-7, atol=1e-9), (k, key)                  # against the CPU reference (file-wide tolerance)
            scale = max(np.abs(c[key]).max(), 1e-300)
            assert np.abs(b[key] - c[key]).max() <= 1e-12 * scale, (k, key, np.abs(b[key] - c[key]).max() / scale)   # same trajectory, two ways out


@needs_binaries
@pytest.mark.parametrize("what,xml", [
    ("planar long-range correction", '<longrange type="planar"><slabs>10</slabs></longrange></algorithm>'),
    ("TemperatureControl thermostat", '<thermostats><thermostat type="TemperatureControl"></thermostat></thermostats></algorithm>')])
def test_host_side_per_molecule_passes_are_rejected_at_start_up(what, xml, tmp_path):
    """features of the reference that loop over host molecules inside the time loop would find empty cells in the device phase:
    the container ends the run while reading the configuration instead of skipping them silently"""
    cfg = M.write_case(str(tmp_path), "ex_Argon", 2, "fused")
    text = open(cfg).read().replace("</algorithm>", xml)
    open(cfg, "w").write(text)
    out = M.run(M.GPU_EXE, cfg, 2, threads=2, check=False)
    assert "GpuLinkedCells: the %s works on host molecules inside the time loop and is not supported" % what in out["log"]
    assert "device context created" not in out["log"]


@needs_binaries
def test_plugins_that_may_change_molecules_end_the_run(tmp_path):
    """a plugin that is not a known read-only writer (COMaligner moves every molecule) would act on a snapshot and silently have no
    effect on the device-resident trajectory: the container ends the run while reading the configuration; a user who knows the
    plugin only reads can say so (<unknownplugins>readonly</unknownplugins>), and then gets a mirror on every step"""
    aligner = '<plugin name="COMaligner"><x>1</x><y>1</y><z>1</z><interval>1</interval><correctionFactor>.5</correctionFactor></plugin>'
    def with_plugin(cfg):      # a simulation plugin: <plugin> directly under <simulation> (Simulation.cpp:520-534)
        text = open(cfg).read().replace("</simulation>", aligner + "</simulation>")
        assert aligner in text
        open(cfg, "w").write(text)
        return cfg
    cfg = with_plugin(M.write_case(str(tmp_path), "ex_Argon", 2, "fused"))
    out = M.run(M.GPU_EXE, cfg, 2, threads=2, check=False)
    assert "GpuLinkedCells: plugin COMaligner is not known to be a read-only end-of-step writer" in out["log"]
    assert "Plugins that change molecules are not supported by the device container" in out["log"]
    assert "device context(s) created" not in out["log"] and "no CUDA device" not in out["log"]
    cfg = with_plugin(M.write_case(str(tmp_path), "ex_Argon", 2, "fused-ro", container_options="<unknownplugins>readonly</unknownplugins>"))
    out = M.run(M.GPU_EXE, cfg, 2, threads=2, check=False)
    assert "(<unknownplugins>readonly</unknownplugins>)" in out["log"]
    assert "GpuLinkedCells: plugin COMaligner reads molecules every 1 step(s)" in out["log"]
    assert "mirror frequency 1 (from the plugins' write frequencies)" in out["log"]


FLOP_PLUGIN = '<outputplugin name="FlopRateWriter"><mode>stdout</mode><writefrequency>2</writefrequency></outputplugin>'


def _flop_lines(log):
    """FlopRateWriter's count line and the lines of FlopCounter::printStats (FlopCounter.cpp:111-122, FlopCounter.h:183-193); the
    FLOP rates depend on the clock and are left out"""
    keep = []
    for line in log.splitlines():
        for key in ("FLOP-Count per Iteration", "Molecule distances:", "kernel calls:", "fraction of Flops for molecule dist"):
            if key in line:
                keep.append(line[line.index(key) - (40 if key == "kernel calls:" else 0):].split("\t")[-1].strip())
    return keep


@pytest.mark.gpu
@needs_binaries
@pytest.mark.parametrize("name,variant", [("ex_Argon", "fused"), ("ex_CO2", "fused"), ("ex_CO2", "operators")])
def test_flop_rate_writer_counts_on_the_device(name, variant, tmp_path):
    """<outputplugin name="FlopRateWriter">: FlopRateWriter::afterForces hands the reference's FlopCounter to traverseCells; the
    container counts molecule pairs, in-cutoff centre pairs per potential and macroscopic pairs on the device (ls1g_flop_count) and
    feeds them to the counter (FlopCounter::addDeviceCounts); the FLOP model and the printing are the reference's.  The counts the
    plugin prints (FLOP-Count per Iteration and every line of FlopCounter::printStats) equal the CPU run's."""
    cpu = M.run(M.REF_EXE, M.write_case(str(tmp_path), name, 6, "cpu", output_plugins=FLOP_PLUGIN), 6)
    gpu = M.run(M.GPU_EXE, M.write_case(str(tmp_path), name, 6, variant, output_plugins=FLOP_PLUGIN), 6)
    assert "3 FLOP sample(s) on the device" in gpu["log"]
    assert "mirror frequency 0 (from the plugins' write frequencies)" in gpu["log"]
    a, b = _flop_lines(cpu["log"]), _flop_lines(gpu["log"])
    assert len(a) >= 3 and sum("FLOP-Count per Iteration" in x for x in a) == 3
    assert a == b


RDF_PLUGIN = ('<outputplugin name="RDF"><writefrequency>3</writefrequency><samplingfrequency>1</samplingfrequency><outputprefix>rdf</outputprefix>'
              '<bins>50</bins><intervallength>0.5</intervallength><angularbins>4</angularbins></outputplugin>')


def _rdf_tables(dirname):
    out = {}
    for f in sorted(os.listdir(dirname)):
        if f.endswith(".rdf") or f.endswith(".ardf"):
            rows = [l.split() for l in open(os.path.join(dirname, f)) if l.strip() and not l.lstrip().startswith("#")]
            out[f] = np.array([[float(x) for x in row] for row in rows])
    return out


@pytest.mark.gpu
@needs_binaries
@pytest.mark.parametrize("variant", ["fused", "operators"])
def test_rdf_plugin_samples_on_the_device(variant, tmp_path):
    """<outputplugin name="RDF"> (molecule, site-site and angular RDF of CO2, sampled every step, written every 3rd) in the reference's
    unmodified loop: RDF::afterForces hands its RDFCellProcessor to traverseCells, the container counts the pairs on the device
    (ls1g_rdf) and adds them to the plugin's counters; the files the plugin writes equal the CPU reference's"""
    cpu_dir, gpu_dir = tmp_path / "cpu", tmp_path / "gpu"
    cpu_dir.mkdir(), gpu_dir.mkdir()
    M.run(M.REF_EXE, M.write_case(str(cpu_dir), "ex_CO2", 6, "cpu", output_plugins=RDF_PLUGIN, equilibration=0), 6)
    gpu = M.run(M.GPU_EXE, M.write_case(str(gpu_dir), "ex_CO2", 6, variant, output_plugins=RDF_PLUGIN, equilibration=0), 6)
    assert "6 RDF and 0 FLOP sample(s) on the device" in gpu["log"]
    assert "mirror frequency 0 (from the plugins' write frequencies)" in gpu["log"]
    a, b = _rdf_tables(str(cpu_dir)), _rdf_tables(str(gpu_dir))
    assert len(a) == 4 and sorted(a) == sorted(b)          # rdf_0-0.{3,6}.rdf and .ardf
    same = total = 0
    for f in a:
        assert a[f].shape == b[f].shape and a[f].shape[0] >= 50, f
        assert np.abs(a[f]).sum() > 0
        # a pair within 1e-8 of a bin border may land in the neighbouring bin on one side
        scale = np.maximum(np.abs(a[f]).max(axis=0), 1e-300)
        assert (np.abs(a[f] - b[f]) / scale).max() < 2e-3, f
        same += int((a[f] == b[f]).sum())
        total += a[f].size
    assert same > 0.99 * total


@pytest.mark.gpu
@needs_binaries
@pytest.mark.parametrize("variant", ["fused", "operators"])
@pytest.mark.parametrize("name", ["ex_Argon", "ex_CO2", "ex_EOX"])
def test_reference_time_loop_on_the_gpu_container(name, variant, tmp_path):
    cpu = M.run(M.REF_EXE, M.write_case(str(tmp_path), name, STEPS, "cpu"), STEPS)
    gpu = M.run(M.GPU_EXE, M.write_case(str(tmp_path), name, STEPS, variant), STEPS)
    assert "GpuLinkedCells: 1 device context(s) created" in gpu["log"]
    assert gpu["gpu"]["steps"] == STEPS and gpu["gpu"]["mirrors"] == 1
    assert cpu["res"].shape == gpu["res"].shape == (STEPS + 1, 10)
    # ResultWriter columns: step time U_pot U_pot_avg p p_avg beta_trans beta_rot c_v N
    for col, what, tol in ((2, "U_pot", 1e-8), (4, "p", 1e-8), (6, "beta_trans", 1e-9), (7, "beta_rot", 1e-9), (9, "N", 0.0)):
        assert rel(gpu["res"][:, col], cpu["res"][:, col]) <= tol, (what, rel(gpu["res"][:, col], cpu["res"][:, col]))
    assert np.array_equal(cpu["final"]["id"], gpu["final"]["id"]) and np.array_equal(cpu["final"]["cid"], gpu["final"]["cid"])
    box = M.load(name)["box"]
    dr = cpu["final"]["r"] - gpu["final"]["r"]
    dr -= box * np.round(dr / box)                     # a molecule on the periodic boundary may be wrapped on one side only
    assert np.max(np.abs(dr)) <= 1e-7 * box[0]
    for k in ("v", "q", "D"):
        assert rel(gpu["final"][k], cpu["final"][k]) <= 1e-6, (k, rel(gpu["final"][k], cpu["final"][k]))


@pytest.mark.gpu
@needs_binaries
@pytest.mark.parametrize("name,blocks,variant", [("ex_Argon", 2, "fused"), ("ex_Argon", 4, "fused"), ("ex_Argon", 8, "fused"), ("ex_Argon", 8, "operators"),
                                                  ("ex_CO2", 2, "fused"), ("ex_CO2", 4, "operators"), ("ex_CO2", 8, "fused")])
def test_reference_time_loop_on_several_device_blocks(name, blocks, variant, tmp_path):
    """<devices>N</devices>: ONE MarDynGpu process whose container cuts the box into DomainDecomposition's block grid
    (DomainDecomposition.cpp:19-41,116-125) with one device context per block (here all on the same GPU; on a multi-GPU box block r
    lives on device r), halo and leaving molecules exchanged between the blocks inside the library.  The reference's unmodified time
    loop and Domain see one container: per-step U_pot, p, betas and the final checkpoint against the CPU reference on the same
    (2 x 2 x 2 replicated) example."""
    steps = 12
    threads = min(os.cpu_count() or 1, 8)
    cpu = M.run(M.REF_EXE, M.write_case(str(tmp_path), name, steps, "cpu", replicate=2), steps, threads=threads)
    gpu = M.run(M.GPU_EXE, M.write_case(str(tmp_path), name, steps, "%s-b%d" % (variant, blocks), replicate=2,
                                        container_options="<devices>%d</devices>" % blocks), steps, threads=threads)
    assert "GpuLinkedCells: %d device context(s) created" % blocks in gpu["log"]
    assert gpu["gpu"]["steps"] == steps and gpu["gpu"]["mirrors"] == 1
    assert cpu["res"].shape == gpu["res"].shape == (steps + 1, 10)
    for col, what, tol in ((2, "U_pot", 1e-8), (4, "p", 1e-8), (6, "beta_trans", 1e-9), (7, "beta_rot", 1e-9), (9, "N", 0.0)):
        assert rel(gpu["res"][:, col], cpu["res"][:, col]) <= tol, (what, rel(gpu["res"][:, col], cpu["res"][:, col]))
    assert np.array_equal(cpu["final"]["id"], gpu["final"]["id"]) and np.array_equal(cpu["final"]["cid"], gpu["final"]["cid"])
    box = 2.0 * M.load(name)["box"]
    dr = cpu["final"]["r"] - gpu["final"]["r"]
    dr -= box * np.round(dr / box)
    assert np.max(np.abs(dr)) <= 1e-7 * box[0]
    for k in ("v", "q", "D"):
        assert rel(gpu["final"][k], cpu["final"][k]) <= 1e-6, (k, rel(gpu["final"][k], cpu["final"][k]))


@pytest.mark.gpu
@needs_binaries
def test_mirror_every_step_changes_nothing(tmp_path):
    """<mirrorfrequency>1: the host cells are refilled from the device after every post-kick (the reference's thermostat then scales the
    host copies, the device applies the same betas at the next pre-kick) - the trajectory must be that of the run without mirrors"""
    a = M.run(M.GPU_EXE, M.write_case(str(tmp_path), "ex_EOX", STEPS, "fused"), STEPS)
    b = M.run(M.GPU_EXE, M.write_case(str(tmp_path), "ex_EOX", STEPS, "fused-mirror", container_options="<mirrorfrequency>1</mirrorfrequency>"), STEPS)
    assert a["gpu"]["mirrors"] == 1 and b["gpu"]["mirrors"] == STEPS
    for col in (2, 4, 6, 7, 9):
        assert rel(b["res"][:, col], a["res"][:, col]) <= 1e-12, col
    for k in ("r", "v", "q", "D"):
        assert rel(b["final"][k], a["final"][k]) <= 1e-10, k


@pytest.mark.gpu
@needs_binaries
def test_reference_time_loop_fcc_one_million(tmp_path):
    """bench.py's workload at 64^3 fcc cells (1 048 576 molecules, the reference's own generator, NVT): the neighbour-list path
    inside the reference's time loop against the CPU reference, 8 steps"""
    steps = 8
    threads = min(os.cpu_count() or 1, 16)
    cpu = M.run(M.REF_EXE, M.write_fcc_case(str(tmp_path), 64, steps, False, "fcc_cpu"), steps, threads=threads)
    gpu = M.run(M.GPU_EXE, M.write_fcc_case(str(tmp_path), 64, steps, True, "fcc_gpu"), steps, threads=threads)
    assert cpu["res"].shape == gpu["res"].shape == (steps + 1, 10) and gpu["res"][0, 9] == 1048576
    for col, what, tol in ((2, "U_pot", 1e-9), (4, "p", 1e-8), (6, "beta_trans", 1e-10), (9, "N", 0.0)):
        assert rel(gpu["res"][:, col], cpu["res"][:, col]) <= tol, (what, rel(gpu["res"][:, col], cpu["res"][:, col]))
