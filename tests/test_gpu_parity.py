"""GPU parity tests: the CUDA path (through the C ABI, include/ls1g.h) against (a) the golden outputs of the
unmodified reference (tests/golden/*.npz) and (b) the CPU oracle on seeded inputs.

Bars (BASELINE.json north_star): cell assignment, halo images and cell-pair enumeration bit-exact; per-molecule
F, M, Vi, upot, virial within 1e-10 relative (double build), stated per case below."""
import numpy as np
import pytest

from oracle import oracle as O
from tests.helpers import golden_cases, load, system_and_grid, vec_rel_err, initial_state
from tests.gpu_helpers import ls1g, container_of, by_id, fcc_lattice, random_quaternions

pytestmark = pytest.mark.gpu

FORCE_CASES = golden_cases("vcp_") + golden_cases("per_") + golden_cases("ex_")
# examples/CO2 (charges +21.2 / -10.6 only 0.38 apart: the site forces cancel to 1e-4 of their magnitude) exists as two dumps of the
# unmodified reference (tests/golden/make_golden_r2.py):
#   ex_CO2        AVX2 build: 1/r^2 and 1/sqrt(r^2) are Newton-Raphson refined estimates (RealVecDouble.h:411-541)
#   ex_CO2_novec  -DVECTOR_INSTRUCTIONS=NONE build: IEEE division and sqrt (SIMD_TYPES.h:79-95)
# The reference's own two builds differ by 3.7e-10 in F on this input.  The CUDA path is held to the NOVEC dump at the stated bar of
# 1e-10 (measured: F 7.9e-11, M 1.2e-11, Vi 8.4e-11, U 1.8e-13; profiles/probe_co2_parity.py) and to the AVX2 dump at 2e-9, which
# only allows for the reference's own NR rounding.  ex_CO2Merker(_novec) - the 3 LJ + quadrupole model of
# examples/surface-tension_LRC/CO2_Merker - has no such cancellation and agrees to 1e-13 with both builds.
FORCE_TOL = {"ex_CO2": 2e-9}


def scalar_close(a, b, tol, floor=0.0):
    return abs(a - b) <= tol * max(abs(b), floor)


@pytest.mark.parametrize("name", golden_cases("cellindex_"))
def test_cell_index_bit_exact(name):
    d = load(name)
    with ls1g.Container(d["box"], float(d["rc"][0]), grid_mode=int(d["mode"][0])) as c:
        cpd, cl = c.get_grid()
        assert cpd == list(d["cells_per_dim"])
        assert np.array_equal(np.array(cl), d["cell_length"])
        assert np.array_equal(c.cell_index_of_points(d["points"]), d["cell_index"])
        assert np.array_equal(c.cell_index_of_points(d["points"], host=True), d["cell_index"])


@pytest.mark.parametrize("name", FORCE_CASES)
def test_cell_pairs_bit_exact(name):
    d = load(name)
    with container_of(d) as c:
        cp = c.cell_pairs()
    cp = cp[np.lexsort((cp[:, 2], cp[:, 1], cp[:, 0]))]
    assert np.array_equal(cp, d["cellpairs"])


def _load_into(c, d, periodic):
    if periodic:
        st = initial_state(d)
        c.add_particles(st["id"], st["cid"], st["r"], st["v"], st["q"], st["L"])
        c.exchange()
        return st
    inorder = np.argsort(d["in_id"])
    c.add_particles(d["in_id"][inorder], d["in_cid"][inorder], d["in_r"][inorder], d["in_v"][inorder],
                    d["in_q"][inorder], d["in_L"][inorder])
    c.update()
    return None


@pytest.mark.parametrize("name", FORCE_CASES)
def test_forces_match_reference(name):
    d = load(name)
    periodic = bool(d["meta"][0])
    tol = FORCE_TOL.get(name, 1e-10)
    with container_of(d) as c:
        _load_into(c, d, periodic)
        upot, virial = c.traverse_cells()
        out = by_id(c.download(("id", "cid", "r", "F", "M", "Vi", "cell")))
        allm = c.download(("id", "r", "cell"), include_halo=True)
    own = d["f_owned"].astype(bool)
    order = np.argsort(d["f_id"][own])
    ref = {k: d["f_" + k][own][order] for k in ("id", "r", "F", "M", "Vi", "cell")}
    assert np.array_equal(out["id"], ref["id"].astype(np.uint64))
    assert np.array_equal(out["r"], ref["r"])
    assert np.array_equal(out["cell"], ref["cell"].astype(np.uint32))          # cell assignment bit-exact
    if periodic:                                                               # halo images bit-exact (id, r, cell)
        mine = np.column_stack([allm["id"].astype(np.float64), allm["r"], allm["cell"].astype(np.float64)])
        refa = np.column_stack([d["f_id"].astype(np.float64), d["f_r"], d["f_cell"].astype(np.float64)])
        mine = mine[np.lexsort(mine.T[::-1])]
        refa = refa[np.lexsort(refa.T[::-1])]
        assert mine.shape == refa.shape and np.array_equal(mine, refa)
    if name == "vcp_F0":
        assert np.abs(out["F"] - ref["F"]).max() < 1e-8
    else:
        assert vec_rel_err(out["F"], ref["F"]) < tol
        assert vec_rel_err(out["M"], ref["M"]) < tol
        assert vec_rel_err(out["Vi"], ref["Vi"]) < tol
    u, w = d["upot_virial"]
    assert scalar_close(upot, u, tol, 1.0 if name == "vcp_U0" else 1e-12)
    assert scalar_close(virial, w, tol, 1e-3 if name == "vcp_F0" else 1e-300)


@pytest.mark.parametrize("name", golden_cases("ex_"))
def test_list_force_kernels_match_reference_at_t0(name):
    """The kernels of the device-resident loop - nl_force_kernel (1CLJ) and nl_force_ms_kernel (multi-site), reading the frozen
    neighbour-list layout - pinned to the REFERENCE's own F, M, Vi, upot, virial at the same bar as the per-step kernels: one
    ls1g_run step with dt = 0 leaves every molecule where the fixture has it, builds the lists and runs the list force pass."""
    d = load(name)
    tol = FORCE_TOL.get(name, 1e-10)
    with container_of(d) as c:
        _load_into(c, d, True)
        c.traverse_cells()
        rec = c.run(1, 0.0)
        stats = c.list_stats()
        out = by_id(c.download(("id", "r", "F", "M", "Vi")))
    if name == "ex_Argon" and stats["rebuilds"] == 0:
        # rc = 5.2 sigma at liquid density: the 9-row tile of a block exceeds the build tile (5120 candidates), the library reports
        # it and stays on the per-step path ("capacities adapt") - nothing of the list kernels ran here
        pytest.skip("ex_Argon is too dense for the list buffers: per-step path")
    assert stats["rebuilds"] == 1, "the step did not run in the list layout"
    own = d["f_owned"].astype(bool)
    order = np.argsort(d["f_id"][own])
    ref = {k: d["f_" + k][own][order] for k in ("id", "r", "F", "M", "Vi")}
    assert np.array_equal(out["id"], ref["id"].astype(np.uint64)) and np.array_equal(out["r"], ref["r"])
    for k in ("F", "M", "Vi"):
        assert vec_rel_err(out[k], ref[k]) < tol, (k, vec_rel_err(out[k], ref[k]))
    u, w = d["upot_virial"]
    assert scalar_close(rec[0, 0], u, tol, 1e-12) and scalar_close(rec[0, 1], w, tol, 1e-300)


@pytest.mark.parametrize("name", FORCE_CASES)
def test_flop_count_matches_reference(name):
    """ls1g_flop_count (the reference's FlopCounter as a device cell processor, adapter/FlopCounter.cpp:196-421) against the
    reference's own total for the same cells (fixture 'flops') and the per-potential counts of the oracle restatement: exact"""
    d = load(name)
    periodic = bool(d["meta"][0])
    with container_of(d) as c:
        _load_into(c, d, periodic)
        fc = c.flop_count()
    assert fc["flops"] == float(d["flops"][0])
    s, g = system_and_grid(d)
    ref = O.flop_count(s, g, d["f_cid"], d["f_r"])
    assert fc["molecule_distances"] == ref["molecule_distances"]
    assert np.array_equal(fc["kernel"], ref["kernel"]) and np.array_equal(fc["macro"], ref["macro"])


def test_known_answers_u0_f0():
    """VectorizedCellProcessorTest.cpp:45-134"""
    d = load("vcp_U0")
    with container_of(d) as c:
        _load_into(c, d, False)
        u, w = c.traverse_cells()
        F = c.download(("F",))["F"]
    assert np.allclose(np.abs(F[:, :2]), 24.0, atol=1e-8) and np.allclose(F[:, 2], 0.0)
    assert abs(u) < 1e-8 and abs(w - 96.0) < 1e-8
    d = load("vcp_F0")
    with container_of(d) as c:
        _load_into(c, d, False)
        u, w = c.traverse_cells()
        F = c.download(("F",))["F"]
    assert np.abs(F).max() < 1e-6 and abs(u + 4.0) < 1e-8 and abs(w) < 1e-6


@pytest.mark.parametrize("name", [n for n in FORCE_CASES if int(load(n)["meta"][2]) > 0])
@pytest.mark.parametrize("fused", [True, False])
def test_time_loop_matches_reference(name, fused):
    """Leapfrog + rigid-body rotation + periodic exchange + forces, NVE.  fused: device-resident ls1g_run;
    otherwise the per-operator calls in Simulation::simulate order."""
    d = load(name)
    steps, dt = int(d["meta"][2]), float(d["dt"][0])
    with container_of(d) as c:
        _load_into(c, d, True)
        c.traverse_cells()
        if fused:
            series = c.run(steps, dt)[:, :4]
        else:
            series = np.zeros((steps, 4))
            for s in range(steps):
                c.event_new_timestep(dt)
                c.exchange()
                series[s, :2] = c.traverse_cells()
                c.delete_outer_particles()
                series[s, 2:] = c.event_forces_calculated(dt)[:2]
        out = by_id(c.download(("id", "r", "v", "q", "L", "F", "M")))
    tol = 1e-9 if name == "ex_CO2" else 1e-11
    assert np.array_equal(out["id"], d["post_id"].astype(np.uint64))
    for k in ("r", "v", "q", "L"):
        ref = d["post_" + k]
        assert np.abs(out[k] - ref).max() <= tol * max(np.abs(ref).max(), 1e-300), k
    assert vec_rel_err(out["F"], d["post_F"]) < 100 * tol
    assert vec_rel_err(out["M"], d["post_M"]) < 100 * tol
    ser = d["series"]
    assert np.all(np.abs(series - ser) <= 100 * tol * np.maximum(np.abs(ser).max(axis=0), 1e-300))


# ------------------------------------------------------------------------------------------------------
# seeded inputs against the oracle
# ------------------------------------------------------------------------------------------------------
def _oracle_forces(sysm, g, box, cid, r, q):
    r2, src, hr = O.exchange(g, box, r)
    all_r = np.concatenate([r2, hr])
    all_cid = np.concatenate([cid, cid[src]])
    all_q = np.concatenate([q, q[src]])
    res = O.forces(sysm, g, all_cid, all_r, all_q)
    n = len(r)
    return {k: (res[k][:n] if isinstance(res[k], np.ndarray) else res[k]) for k in res}


@pytest.mark.parametrize("kernel", ["default", "simple"])
def test_1clj_liquid_vs_oracle(kernel, monkeypatch):
    """cfg B in small: perturbed fcc 1CLJ, rho*=0.8442, rc=2.5 (13 molecules per cell): the specialised 1CLJ kernels
    (default: warp-per-cell with FP32 candidate filter; simple: v1 thread-per-molecule)"""
    monkeypatch.setenv("LS1G_FORCE_KERNEL", kernel)
    r, v, box = fcc_lattice(12, seed=3)
    rng = np.random.default_rng(5)
    r = np.mod(r + rng.normal(scale=0.08, size=r.shape), box)
    n = len(r)
    sysm = O.System.single_lj(rc=2.5)
    g = O.grid([0, 0, 0], box, 2.5, 0)
    cid = np.zeros(n, dtype=np.uint32)
    q = np.tile([1.0, 0, 0, 0], (n, 1))
    ref = _oracle_forces(sysm, g, box, cid, r, q)
    with ls1g.Container(box, 2.5) as c:
        c.set_components([ls1g.lj_component()])
        c.add_particles(np.arange(n), cid, r, v)
        c.exchange()
        u, w = c.traverse_cells()
        out = by_id(c.download(("id", "F", "Vi")))
        # molecule pairs inside the cutoff, each once (minimum image, brute force)
        n_in = 0
        for a in range(0, n, 512):
            dd = r[a:a + 512, None, :] - r[None, :, :]
            dd -= box * np.round(dd / box)
            n_in += int(((dd * dd).sum(axis=2) < 2.5 * 2.5).sum()) - len(dd)
        assert c.last_pair_count() == n_in // 2
    assert vec_rel_err(out["F"], ref["F"]) < 1e-10
    assert vec_rel_err(out["Vi"], ref["Vi"]) < 1e-10
    assert scalar_close(u, ref["upot"], 1e-10) and scalar_close(w, ref["virial"], 1e-10)


def test_mixture_vs_oracle():
    """two components (2CLJ+quadrupole and 3LJ+dipole+charges) with xi/eta mixing, rc_LJ < rc, random orientations"""
    rng = np.random.default_rng(11)
    box = np.array([21.0, 17.5, 19.0])
    n = 1500
    # jittered grid so that no two molecules overlap
    m = int(np.ceil(n ** (1 / 3)))
    gi = np.stack(np.meshgrid(*[np.arange(m)] * 3, indexing="ij"), axis=-1).reshape(-1, 3)[:n]
    r = (gi + 0.5 + rng.uniform(-0.2, 0.2, size=(n, 3))) * (box / m)
    cid = rng.integers(0, 2, size=n).astype(np.uint32)
    q = random_quaternions(n, rng)
    counts = [[2, 0, 0, 1, 2, 0], [3, 2, 1, 0, 3, 2]]
    scal = [[2.0, 0.5, 0.5, 0.0], [3.0, 0.4, 0.6, 0.9]]
    lj = [[0, 0, -0.3, 1, 1.0, 1.0, 0.0], [0, 0, 0.3, 1, 1.0, 1.0, 0.0],
          [0.2, 0, 0, 1, 0.8, 0.9, 0.0], [-0.2, 0.1, 0, 1, 1.2, 1.1, 0.0], [0, -0.2, 0.2, 1, 0.5, 0.8, 0.0]]
    qu = [[0, 0, 0, 0, 0, 1, 1.5]]
    ch = [[0.25, 0, 0, 0, 0.4], [-0.25, 0, 0, 0, -0.4]]
    di = [[0, 0.1, 0, 1, 0, 0, 0.7]]
    mix = [0.95, 1.03]
    sysm = O.System(counts, scal, lj, ch, di, qu, mix, 1e10, 3.2, 2.6)
    g = O.grid([0, 0, 0], box, 3.2, 0)
    ref = _oracle_forces(sysm, g, box, cid, r, q)
    with ls1g.Container(box, 3.2, lj_cutoff=2.6) as c:
        c.set_components(ls1g.components_from_tables(counts, scal, lj, ch, di, qu), mix)
        c.add_particles(np.arange(n), cid, r, None, q, None)
        c.exchange()
        u, w = c.traverse_cells()
        out = by_id(c.download(("id", "F", "M", "Vi")))
    assert vec_rel_err(out["F"], ref["F"]) < 1e-10
    assert vec_rel_err(out["M"], ref["M"]) < 1e-10
    assert vec_rel_err(out["Vi"], ref["Vi"]) < 1e-10
    assert scalar_close(u, ref["upot"], 1e-10) and scalar_close(w, ref["virial"], 1e-10)


def test_empty_and_single_molecule():
    """edge cases: empty container, one molecule (no partner), ragged cells"""
    with ls1g.Container([10.0, 10.0, 10.0], 2.5) as c:
        c.set_components([ls1g.lj_component()])
        c.add_particles(np.zeros(0, dtype=np.uint64), np.zeros(0, dtype=np.uint32), np.zeros((0, 3)))
        c.exchange()
        assert c.traverse_cells() == (0.0, 0.0)
        assert c.count() == (0, 0)
        c.add_particles([7], [0], [[0.1, 9.9, 5.0]], [[1.0, 0.0, 0.0]])
        c.exchange()
        assert c.count() == (1, 3)          # images across the x and y faces and the xy edge
        u, w = c.traverse_cells()
        assert u == 0.0 and w == 0.0
        rec = c.run(3, 0.5)
        out = c.download(("id", "r", "v", "F"))
        assert np.array_equal(out["F"], np.zeros((1, 3)))
        assert np.allclose(out["r"], [[1.6, 9.9, 5.0]]) and rec[-1, 4] == 1


def test_leaving_molecules_wrap():
    """molecules moved outside the box re-enter on the other side (handleDomainLeavingParticles), ids conserved"""
    r, v, box = fcc_lattice(6, seed=2)
    n = len(r)
    sysm = O.System.single_lj(rc=2.5)
    g = O.grid([0, 0, 0], box, 2.5, 0)
    rng = np.random.default_rng(4)
    r = r + rng.uniform(-1.0, 1.0, size=r.shape)       # some leave through faces, edges and corners
    r_wrapped, src, hr = O.exchange(g, box, r)
    with ls1g.Container(box, 2.5) as c:
        c.set_components([ls1g.lj_component()])
        c.add_particles(np.arange(n), np.zeros(n, dtype=np.uint32), r, v)
        c.exchange()
        out = by_id(c.download(("id", "r", "cell")))
        allm = c.download(("id", "r"), include_halo=True)
    assert np.array_equal(out["r"], r_wrapped)
    assert np.array_equal(out["cell"], O.cell_indices(g, r_wrapped))
    mine = np.column_stack([allm["id"].astype(np.float64), allm["r"]])
    ref = np.column_stack([np.concatenate([np.arange(n), src]).astype(np.float64), np.concatenate([r_wrapped, hr])])
    assert np.array_equal(mine[np.lexsort(mine.T[::-1])], ref[np.lexsort(ref.T[::-1])])


def test_update_then_exchange_keeps_the_leavers():
    """The reference's order in Simulation::updateParticleContainerAndDecomposition (Simulation.cpp:1193-1215): LinkedCells::update
    first (leavers move into halo cells and stay molecules), then balanceAndExchange wraps them.  ls1g_update must not turn an
    owned leaver into a halo copy that the next exchange drops."""
    r, v, box = fcc_lattice(6, seed=5)
    n = len(r)
    g = O.grid([0, 0, 0], box, 2.5, 0)
    v = v * 40.0                                           # one pre-kick/drift carries molecules near the faces out of the box
    with ls1g.Container(box, 2.5) as c:
        c.set_components([ls1g.lj_component()])
        c.add_particles(np.arange(n), np.zeros(n, dtype=np.uint32), r, v)
        c.exchange()
        c.traverse_cells()
        c.delete_outer_particles()
        c.event_new_timestep(0.02)
        moved = by_id(c.download(("id", "r")))
        assert np.any(moved["r"] < 0) or np.any(moved["r"] >= box)
        c.update()
        assert c.count()[0] == n
        after_update = by_id(c.download(("id", "r", "cell")))
        assert np.array_equal(after_update["id"], np.arange(n)) and np.array_equal(after_update["r"], moved["r"])
        assert np.array_equal(after_update["cell"], O.cell_indices(g, moved["r"]))
        c.exchange()
        assert c.count()[0] == n
        out = by_id(c.download(("id", "r", "cell")))
    r_wrapped, _, _ = O.exchange(g, box, moved["r"])
    assert np.array_equal(out["id"], np.arange(n))
    assert np.array_equal(out["r"], r_wrapped)
    assert np.array_equal(out["cell"], O.cell_indices(g, r_wrapped))


# ------------------------------------------------------------------------------------------------------
# size-independent properties at a larger size
# ------------------------------------------------------------------------------------------------------
def test_nve_properties_large():
    """256k-molecule 1CLJ fcc liquid: Newton 3 (sum F = 0), momentum and molecule count conserved, NVE energy
    drift small over 200 steps, re-sort idempotent, deterministic (bitwise identical repeat).  The potential is the
    SHIFTED LJ here (Component::addLJcenter shift6, Component.cpp:105-117): the unshifted one jumps by U(rc) whenever a
    pair crosses the cutoff, so its "energy" is not a conserved quantity (see test_nve_drift_matches_oracle)."""
    r, v, box = fcc_lattice(40, seed=9)
    n = len(r)
    dt = 0.002
    s6 = (1.0 / 2.5) ** 6
    shift6 = 24.0 * (s6 - s6 * s6)

    def one():
        with ls1g.Container(box, 2.5) as c:
            c.set_components([ls1g.lj_component(shift6=shift6)])
            c.add_particles(np.arange(n), np.zeros(n, dtype=np.uint32), r, v)
            c.exchange()
            u0, _ = c.traverse_cells()
            first = c.download(("id", "r", "F"))
            c.exchange()                                   # idempotent: same order, same content
            again = c.download(("id", "r"))
            assert np.array_equal(first["id"], again["id"]) and np.array_equal(first["r"], again["r"])
            c.traverse_cells()
            rec = c.run(200, dt)
            return u0, first, rec, by_id(c.download(("id", "r", "v", "F")))

    u0, first, rec, out = one()
    fmax = np.abs(first["F"]).max()
    assert fmax == 0.0 or np.abs(first["F"].sum(axis=0)).max() < 1e-9 * max(fmax, 1.0) * np.sqrt(n)
    assert np.all(rec[:, 4] == n)
    e = rec[:, 0] + 0.5 * rec[:, 2]                        # Upot + Ekin (summv2 = sum m v^2)
    assert np.abs(e - e[0]).max() < 1e-4 * abs(e[0])       # leapfrog, dt*=0.002, shifted potential
    assert np.abs(out["v"].sum(axis=0)).max() < 1e-8 * n
    assert np.abs(out["F"].sum(axis=0)).max() < 1e-9 * np.abs(out["F"]).max() * np.sqrt(n)
    u0b, _, recb, outb = one()
    assert u0 == u0b and np.array_equal(rec, recb) and np.array_equal(out["r"], outb["r"])


def test_nve_drift_matches_oracle():
    """NVE over 300 steps (unshifted LJ, as cfg B): the energy series of the device loop follows the oracle's, so the
    drift is the reference algorithm's own (BASELINE.json: drift no worse than the reference)"""
    r, v, box = fcc_lattice(8, seed=4)
    n = len(r)
    steps, dt = 300, 0.002
    sysm = O.System.single_lj(rc=2.5)
    g = O.grid([0, 0, 0], box, 2.5, 0)
    cid = np.zeros(n, dtype=np.uint32)
    q = np.tile([1.0, 0, 0, 0], (n, 1))
    f0 = _oracle_forces(sysm, g, box, cid, r, q)
    ref = O.run(sysm, g, box, cid, r, v, q, np.zeros((n, 3)), f0["F"], np.zeros((n, 3)), steps, dt)
    with ls1g.Container(box, 2.5) as c:
        c.set_components([ls1g.lj_component()])
        c.add_particles(np.arange(n), cid, r, v)
        c.exchange()
        c.traverse_cells()
        rec = c.run(steps, dt)
    e_ref = ref["series"][:, 0] + 0.5 * ref["series"][:, 2]
    e = rec[:, 0] + 0.5 * rec[:, 2]
    assert np.abs(e - e_ref).max() < 1e-8 * abs(e_ref[0])
    assert abs((e[-1] - e[0]) - (e_ref[-1] - e_ref[0])) < 1e-8 * abs(e_ref[0])


@pytest.mark.parametrize("which", ["unshifted", "shifted"])
def test_nve_1000_steps_against_the_reference_record(which):
    """BASELINE.json: "NVE energy drift must be no worse than the reference over 1000 steps".  tests/golden/ref_nve_1000_steps.json is the
    UNMODIFIED reference (LinkedCells + C08 + VectorizedCellProcessor AVX2 + Leapfrog, thermostat off) stepping 4 * 64^3 = 1 048 576
    molecules of the bench workload for 1000 steps (tests/golden/make_golden_r2.py nve, ~12 min of CPU each).  The device loop on the same
    input: total energy per molecule at every 100th step equal to the reference's to 1e-6 (the trajectories are chaotic, the energy is
    not), drift no worse than the reference's, same temperature."""
    import json
    import os
    import importlib
    ref_all = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_nve_1000_steps.json")))
    ref = ref_all[which]
    gen = importlib.import_module("ls1-mardyn_b200.generators")
    ids, r, v = gen.fcc_block(64, rho=0.8442, temperature=1.44, seed=1)
    _, L = gen.fcc_box(64, 0.8442)
    v -= v.mean(axis=0)
    n = len(ids)
    assert n == ref_all["n_molecules"] and ref_all["steps"] == 1000
    with ls1g.Container(list(L), 2.5) as c:
        c.set_components([ls1g.lj_component(shift6=ref["shift6"])])
        c.add_particles(ids, np.zeros(n, dtype=np.uint32), r, v)
        c.exchange()
        c.traverse_cells()
        rec = c.run(1000, 0.002)
    e = (rec[:, 0] + 0.5 * rec[:, 2]) / n
    e_ref = np.array(ref["e_total_per_molecule_every_100"])
    assert np.abs(e[::100] - e_ref).max() <= 1e-6 * abs(e_ref[0]), np.abs(e[::100] - e_ref).max()
    drift = (e[-1] - e[0]) / abs(e[0])
    assert abs(drift) <= abs(ref["rel_energy_drift"]) * 1.01 + 1e-9, (drift, ref["rel_energy_drift"])
    assert abs(drift - ref["rel_energy_drift"]) <= 1e-6
    T = rec[:, 2] / (3.0 * n)
    assert abs(T[0] - ref["temperature_first_last"][0]) < 1e-9 and abs(T[-1] - ref["temperature_first_last"][1]) < 1e-4


def test_list_loop_equals_per_step_loop(monkeypatch):
    """ls1g_run with neighbour lists (rebuild on displacement > skin/2, frozen layout, refreshed images) against the
    rebuild-every-step loop (LS1G_SKIN=0) on a liquid that crosses the periodic boundary: same pairs, same energies"""
    r, v, box = fcc_lattice(14, seed=12, temperature=2.0)
    n = len(r)
    out = {}
    # "0.5+0.1": outer lists with skin 0.5, pruned every few steps to inner lists with skin 0.1 (LS1G_SKIN_INNER, lists.cuh: nl_prune_kernel)
    for skin in ("0", "0.3", "0.6", "0.5+0.1"):
        monkeypatch.setenv("LS1G_SKIN", skin.split("+")[0])
        monkeypatch.setenv("LS1G_SKIN_INNER", skin.split("+")[1] if "+" in skin else "0")
        with ls1g.Container(box, 2.5) as c:
            c.set_components([ls1g.lj_component()])
            c.add_particles(np.arange(n), np.zeros(n, dtype=np.uint32), r, v)
            c.exchange()
            c.traverse_cells()
            rec = c.run(120, 0.004)
            pairs = c.last_pair_count()
            st = by_id(c.download(("id", "r", "v", "F", "Vi", "cell")))
            # the container is left in the per-step state: a plain exchange + traverse reproduces the forces
            c.exchange()
            u, w = c.traverse_cells()
            st2 = by_id(c.download(("id", "F")))
            assert vec_rel_err(st2["F"], st["F"]) < 1e-12
        out[skin] = (rec, st, pairs)
    ref_rec, ref_st, ref_pairs = out["0"]
    assert np.all(ref_st["r"] >= 0) and np.all(ref_st["r"] < box)
    for skin in ("0.3", "0.6", "0.5+0.1"):
        rec, st, pairs = out[skin]
        assert pairs == ref_pairs
        assert np.all(st["r"] >= 0) and np.all(st["r"] < box)
        assert np.array_equal(st["cell"], ref_st["cell"])
        assert np.all(np.abs(rec - ref_rec) <= 1e-9 * np.maximum(np.abs(ref_rec).max(axis=0), 1e-300))
        for k in ("r", "v"):
            assert np.abs(st[k] - ref_st[k]).max() <= 1e-9 * np.abs(ref_st[k]).max(), k
        assert vec_rel_err(st["F"], ref_st["F"]) < 1e-7
        assert vec_rel_err(st["Vi"], ref_st["Vi"]) < 1e-7


@pytest.mark.parametrize("name,list_mode", [("ex_EOX", False), ("ex_CO2", False), ("ex_Argon", True)])
def test_selective_thermostat_matches_oracle(name, list_mode):
    """Domain::calculateGlobalValues explosion heuristics (Domain.cpp:254-295): per-molecule caps on U_trans (v, F) and U_rot (D, M),
    against the numpy restatement; limits chosen so that roughly half of the molecules are capped.  list_mode: applied while the
    container is in the frozen neighbour-list layout of ls1g_run (velocities only, no re-sort)."""
    d = load(name)
    st = initial_state(d)
    sysm, _ = system_and_grid(d)
    rng = np.random.default_rng(11)
    st["L"] = st["L"] + 0.02 * rng.normal(size=st["L"].shape) * (d["comp_scalars"][st["cid"], 1:4] != 0)
    with container_of(d) as c:
        c.add_particles(st["id"], st["cid"], st["r"], st["v"], st["q"], st["L"])
        c.exchange()
        c.traverse_cells()
        c.delete_outer_particles()
        if list_mode:
            c.run(2, float(d["dt"][0]))
        fields = ("id", "cid", "v", "q", "L", "F", "M")
        m = d["comp_scalars"][0, 0]
        if not list_mode:
            before = by_id(c.download(fields))
        else:
            # a download would make c leave the list layout: read the same state from a twin container that took the same two steps
            with container_of(d) as c2:
                c2.add_particles(st["id"], st["cid"], st["r"], st["v"], st["q"], st["L"])
                c2.exchange(); c2.traverse_cells(); c2.delete_outer_particles()
                c2.run(2, float(d["dt"][0]))
                before = by_id(c2.download(fields))
        ut = 0.5 * m * (before["v"] ** 2).sum(axis=1)
        limit = float(np.median(ut))
        Ti = limit / 10.0
        c.selective_thermostat(limit, Ti)
        after = by_id(c.download(fields))
    v, F, L, M = O.selective_thermostat(sysm, before["cid"], before["v"], before["F"], before["q"], before["L"], before["M"], limit, Ti)
    capped = 0.5 * m * (after["v"] ** 2).sum(axis=1)
    assert np.all(capped <= limit * (1 + 1e-12)) and 0.3 < np.mean(ut > limit) < 0.7
    tol = 1e-7 if list_mode else 1e-13
    for k, ref in (("v", v), ("F", F), ("L", L), ("M", M)):
        scale = max(np.abs(ref).max(), 1e-300)
        assert np.abs(after[k] - ref).max() <= tol * scale, (k, np.abs(after[k] - ref).max() / scale)


@pytest.mark.parametrize("name", ["ex_Argon", "ex_EOX"])
def test_host_driven_step_equals_scale_plus_run(name):
    """ls1g_step(dt, betas) = ls1g_scale(betas) + ls1g_run(1): the entry point the reference's time loop drives through
    GpuLinkedCells (thermostat of the previous step fused into the pre-kick); Argon takes the neighbour-list path, EOX the per-step one"""
    d = load(name)
    st = initial_state(d)
    dt = float(d["dt"][0])
    betas = [(1.0, 1.0), (0.97, 0.99), (1.02, 1.0), (0.995, 1.01), (1.0, 1.0), (0.9, 0.95)]
    out = []
    for fused in (False, True):
        with container_of(d) as c:
            c.add_particles(st["id"], st["cid"], st["r"], st["v"], st["q"], st["L"])
            c.exchange()
            c.traverse_cells()
            c.delete_outer_particles()
            recs = []
            for k, (bt, br) in enumerate(betas):
                if fused:
                    recs.append(c.step(dt, bt, br, vi=(k == len(betas) - 1)))
                else:
                    c.scale(bt, br)
                    recs.append(c.run(1, dt)[0])
            out.append((np.array(recs), by_id(c.download(("id", "r", "v", "q", "L", "F", "M", "Vi")))))
    (ra, sa), (rb, sb) = out
    assert np.all(np.abs(ra - rb) <= 1e-11 * np.maximum(np.abs(ra), 1e-300)), np.abs(ra - rb).max()
    for k in ("r", "v", "q", "L", "F", "M", "Vi"):
        scale = max(np.abs(sa[k]).max(), 1e-300)
        assert np.abs(sa[k] - sb[k]).max() <= 1e-10 * scale, (k, np.abs(sa[k] - sb[k]).max() / scale)


def test_thermostat_run_holds_temperature():
    """NVT: Domain::calculateGlobalValues betas (Domain.cpp:226-247) drive T to the target"""
    r, v, box = fcc_lattice(10, seed=1, temperature=1.0)
    n = len(r)
    with ls1g.Container(box, 2.5) as c:
        c.set_components([ls1g.lj_component()])
        c.add_particles(np.arange(n), np.zeros(n, dtype=np.uint32), r, v)
        c.exchange()
        c.traverse_cells()
        rec = c.run(50, 0.002, target_temperature=1.3)
        T = rec[:, 2] / (3 * rec[:, 4])
        assert abs(T[-1] - 1.3) < 0.05
        # oracle with the same thermostat rule
        sysm = O.System.single_lj(rc=2.5)
        g = O.grid([0, 0, 0], box, 2.5, 0)
        q = np.tile([1.0, 0, 0, 0], (n, 1))
        f0 = _oracle_forces(sysm, g, box, np.zeros(n, dtype=np.uint32), r, q)
        ref = O.run(sysm, g, box, np.zeros(n, dtype=np.uint32), r, v, q, np.zeros((n, 3)), f0["F"], np.zeros((n, 3)),
                    50, 0.002, 1.3)
        assert np.all(np.abs(rec[:, :3] - ref["series"][:, :3]) <= 1e-9 * np.abs(ref["series"][:, :3]).max(axis=0))


# ------------------------------------------------------------------------------------------------------
# block decomposition: several ranks as contexts of this process (same pack/unpack kernels as the NCCL path)
# ------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n_ranks", [2, 4, 8])
@pytest.mark.parametrize("case", ["1clj", "ex_CO2"])
def test_decomposed_matches_single_rank(n_ranks, case):
    if case == "1clj":
        r, v, box = fcc_lattice(12, seed=6)
        rng = np.random.default_rng(8)
        r = np.mod(r + rng.normal(scale=0.05, size=r.shape), box)
        n = len(r)
        ids, cid, q, L = np.arange(n), np.zeros(n, dtype=np.uint32), None, None
        rc, rclj, comps, mix, dt = 2.5, 2.5, [ls1g.lj_component()], (), 0.004
    else:
        d = load(case)
        st = initial_state(d)
        ids, cid, r, v, q, L = st["id"], st["cid"], st["r"], st["v"], st["q"], st["L"]
        box, rc, rclj, dt = d["box"], float(d["rc"][0]), float(d["rc"][1]), float(d["dt"][0])
        from tests.gpu_helpers import components_of
        comps, mix = components_of(d), d["mixcoeff"]
        n = len(ids)
    steps = 4
    with ls1g.Container(box, rc, lj_cutoff=rclj) as c:
        c.set_components(comps, mix)
        c.add_particles(ids, cid, r, v, q, L)
        c.exchange()
        c.traverse_cells()
        ref_rec = c.run(steps, dt)
        ref = by_id(c.download(("id", "r", "v", "F", "M", "q", "L")))
    ranks = []
    try:
        for rank in range(n_ranks):
            grid, coords = ls1g.decomp_grid(n_ranks, rank)
            bmin, bmax = ls1g.decomp_box(grid, coords, box)
            c = ls1g.Container(bmax, rc, box_min=bmin, lj_cutoff=rclj, global_length=box, rank_grid=grid,
                               rank_coords=coords, capacity=4 * n + 1024)
            c.set_components(comps, mix)
            mine = np.all((r >= np.array(bmin)) & (r < np.array(bmax)), axis=1)
            c.add_particles(ids[mine], cid[mine], r[mine], v[mine], None if q is None else q[mine], None if L is None else L[mine])
            ranks.append(c)
        ls1g.comm_init_local(ranks)
        ls1g.exchange_group(ranks)
        rec = np.zeros((steps, 6))
        for c in ranks:
            c.traverse_cells()
        for s in range(steps):
            for c in ranks:
                c.event_new_timestep(dt)
            ls1g.exchange_group(ranks)
            for c in ranks:
                u, w = c.traverse_cells()
                c.delete_outer_particles()
                a, b, nn, rd = c.event_forces_calculated(dt)
                rec[s] += [u, w, a, b, nn, rd]
        parts = [c.download(("id", "r", "v", "F", "M", "q", "L")) for c in ranks]
    finally:
        for c in ranks:
            c.close()
    out = by_id({k: np.concatenate([p[k] for p in parts]) for k in parts[0]})
    assert np.array_equal(out["id"], ref["id"])
    assert np.all(rec[:, 4] == n)
    tol = 1e-9 if case == "ex_CO2" else 1e-11
    for k in ("r", "v", "q", "L"):
        assert np.abs(out[k] - ref[k]).max() <= tol * max(np.abs(ref[k]).max(), 1e-300), k
    assert vec_rel_err(out["F"], ref["F"]) < 100 * tol
    assert np.all(np.abs(rec[:, :4] - ref_rec[:, :4]) <= 100 * tol * np.maximum(np.abs(ref_rec[:, :4]).max(axis=0), 1e-300))


def _group_of_ranks(n_ranks, box, rc, ids, r, v, capacity):
    ranks = []
    for rank in range(n_ranks):
        grid, coords = ls1g.decomp_grid(n_ranks, rank)
        bmin, bmax = ls1g.decomp_box(grid, coords, box)
        # all ranks on this GPU on a one-GPU box; spread over the GPUs (peer copies over NVLink) where there are several
        c = ls1g.Container(bmax, rc, box_min=bmin, global_length=box, rank_grid=grid, rank_coords=coords, capacity=capacity,
                           device=rank % ls1g.device_count())
        c.set_components([ls1g.lj_component()])
        mine = np.all((r >= np.array(bmin)) & (r < np.array(bmax)), axis=1)
        c.add_particles(ids[mine], np.zeros(int(mine.sum()), dtype=np.uint32), r[mine], v[mine])
        ranks.append(c)
    ls1g.comm_init_local(ranks)
    ls1g.exchange_group(ranks)
    for c in ranks:
        c.traverse_cells()
    return ranks


def _run_group_case(n_ranks, box, rc, ids, r, v, steps, dt):
    ranks = _group_of_ranks(n_ranks, box, rc, ids, r, v, capacity=4 * len(ids) + 4096)
    try:
        rec = ls1g.run_group(ranks, steps, dt)
        stats = ranks[0].list_stats()
        parts = [c.download(("id", "r", "v", "F")) for c in ranks]      # leaves the list layout: one exchange of the whole group
    finally:
        for c in ranks:
            c.close()
    return rec, by_id({k: np.concatenate([p[k] for p in parts]) for k in parts[0]}), stats


@pytest.mark.parametrize("n_ranks", [2, 4, 8])
def test_group_list_loop_matches_oracle(n_ranks, monkeypatch):
    """The code path the multi-GPU bench times - neighbour lists over a block decomposition: two-pass build exchange, persistent
    halo lists refreshed every step, two-phase force launch overlapped with the refresh, guarded speculative steps, rebuilds agreed on
    the all-reduced displacement - driven here through the in-process transport (all ranks are contexts on this GPU; only the
    ncclSend/ncclRecv calls are replaced by event-ordered copies).  Held to the CPU oracle's trajectory of the same global system
    (per-step upot, virial, kinetic sums; final r, v, F), 72 steps with several rebuilds; overlap on and off are bitwise equal."""
    r, v, box = fcc_lattice(12, seed=6, temperature=2.0)
    rng = np.random.default_rng(8)
    r = np.mod(r + rng.normal(scale=0.05, size=r.shape), box)
    n = len(r)
    ids, cid = np.arange(n), np.zeros(n, dtype=np.uint32)
    steps, dt, rc = 72, 0.004, 2.5
    sysm = O.System.single_lj(rc=rc)
    g = O.grid([0, 0, 0], box, rc, 0)
    q = np.tile([1.0, 0, 0, 0], (n, 1))
    f0 = _oracle_forces(sysm, g, box, cid, r, q)
    ref = O.run(sysm, g, box, cid, r, v, q, np.zeros((n, 3)), f0["F"], np.zeros((n, 3)), steps, dt)
    out = {}
    for overlap in ("2", "1", "0"):      # boundary blocks next to the inner ones (default) / after them on the main stream / no overlap
        monkeypatch.setenv("LS1G_OVERLAP", overlap)
        out[overlap] = _run_group_case(n_ranks, box, rc, ids, r, v, steps, dt)
    rec, st, stats = out["1"]
    assert stats["rebuilds"] >= 3, stats                              # the first build and at least two displacement-triggered ones
    assert np.all(rec[:, 4] == n)
    ser = ref["series"]
    assert np.all(np.abs(rec[:, :3] - ser[:, :3]) <= 1e-9 * np.abs(ser[:, :3]).max(axis=0)), np.abs(rec[:, :3] - ser[:, :3]).max(axis=0)
    order = np.argsort(ids)
    dr = np.abs(st["r"] - ref["r"][order])
    dr = np.minimum(dr, np.abs(dr - np.asarray(box)))                 # a molecule exactly on the periodic boundary may sit on either side
    assert dr.max() < 1e-9 and np.abs(st["v"] - ref["v"][order]).max() < 1e-8
    assert vec_rel_err(st["F"], ref["F"][order]) < 1e-6                # trajectory divergence over 72 steps, not arithmetic
    for other in ("0", "2"):      # every block owns its outputs and partial sums: the three launch schemes are bitwise equal
        rec0, st0, _ = out[other]
        assert np.array_equal(rec, rec0), other
        for k in ("id", "r", "v", "F"):
            assert np.array_equal(st[k], st0[k]), (other, k)


@pytest.mark.parametrize("skin", ["0.04", "0.11", "0.3+0.05"])
def test_group_list_loop_with_lists_expiring_in_flight(skin, monkeypatch):
    """Lists that expire every two to four steps while up to three guarded steps are already enqueued: the skipped steps (odd and even
    counts; the fused kernel swaps the position buffers on every launch, skipped or not) are repeated after the rebuild.  Same pairs
    and the same trajectory as the rebuild-every-step loop of the same group (LS1G_SKIN=0), and as a single rank."""
    r, v, box = fcc_lattice(12, seed=16, temperature=3.0)
    n = len(r)
    ids = np.arange(n)
    steps, dt, rc = 40, 0.004, 2.5
    res = {}
    for sk in ("0", skin):      # "a+b": outer skin a, inner lists with skin b that expire every other step (prunes enqueued ahead, refused, repeated)
        monkeypatch.setenv("LS1G_SKIN", sk.split("+")[0])
        monkeypatch.setenv("LS1G_SKIN_INNER", sk.split("+")[1] if "+" in sk else "0")
        res[sk] = _run_group_case(4, box, rc, ids, r, v, steps, dt)
    with ls1g.Container(box, rc) as c:                                  # single rank, lists with the same small skin
        c.set_components([ls1g.lj_component()])
        c.add_particles(ids, np.zeros(n, dtype=np.uint32), r, v)
        c.exchange()
        c.traverse_cells()
        rec1 = c.run(steps, dt)
        st1 = by_id(c.download(("id", "r", "v", "F")))
        single_stats = c.list_stats()
    rec, st, stats = res[skin]
    rec0, st0, stats0 = res["0"]
    least = 2 if "+" in skin else steps // 5
    assert stats0["rebuilds"] == 0 and stats["rebuilds"] >= least and single_stats["rebuilds"] >= least, (stats0, stats, single_stats)
    for other_rec, other_st in ((rec0, st0), (rec1, st1)):
        assert np.all(np.abs(rec - other_rec) <= 1e-9 * np.maximum(np.abs(other_rec).max(axis=0), 1e-300))
        for k in ("r", "v"):
            assert np.abs(st[k] - other_st[k]).max() <= 1e-9 * np.abs(other_st[k]).max(), k
        assert np.array_equal(st["id"], other_st["id"])


@pytest.mark.parametrize("name,dt_factor,steps,min_rebuilds", [("ex_CO2", 4.0, 60, 3), ("ex_EOX", 1.0, 60, 3), ("ex_CO2Merker", 1.0, 60, 3),
                                                              ("mixture", 4.0, 40, 2)])
def test_multisite_list_loop_equals_per_step_loop(name, dt_factor, steps, min_rebuilds, monkeypatch):
    """Rigid multi-site molecules on the neighbour lists (centre-centre lists, site loops inside; orientations of the periodic images
    refreshed with their source) against the rebuild-every-step cell scan (LS1G_LIST_MULTISITE=0): same pair count, same trajectory.
    Step and length are chosen per input so that the lists expire several times while the (unequilibrated) inputs stay intact:
    EOX at 600 K and CO2-Merker rebuild every ~5 steps at their own step.  "mixture": the CO2 example with every second molecule
    turned into a second component (same geometry, LJ epsilon x 0.8, sigma x 1.03; xi = 0.9, eta = 1.05), which takes the kernels
    with run-time site counts and the LJ tables of all component pairs on both paths.  (The reference's own multi-component
    vectorisation input cannot be time-stepped: it contains components of dipole/quadrupole sites only, i.e. of zero mass.)"""
    from tests.gpu_helpers import components_of
    d = load("ex_CO2" if name == "mixture" else name)
    st = initial_state(d)
    dt = dt_factor * float(d["dt"][0])
    comps, mix, cid = components_of(d), d["mixcoeff"], st["cid"]
    if name == "mixture":
        second = {k: (np.array(v, dtype=np.float64).copy() if isinstance(v, np.ndarray) else v) for k, v in comps[0].items()}
        second["lj"][:, 4] *= 0.8
        second["lj"][:, 5] *= 1.03
        comps, mix = [comps[0], second], [0.9, 1.05]
        cid = (st["id"] % 2).astype(np.uint32)
    out = {}
    for mode in ("1", "0"):
        monkeypatch.setenv("LS1G_LIST_MULTISITE", mode)
        with ls1g.Container(d["box"], float(d["rc"][0]), lj_cutoff=float(d["rc"][1]), eps_rf=float(d["epsRF"][0]), grid_mode=int(d["meta"][1])) as c:
            c.set_components(comps, mix)
            c.add_particles(st["id"], cid, st["r"], st["v"], st["q"], st["L"])
            c.exchange()
            c.traverse_cells()
            rec = c.run(steps, dt)
            out[mode] = (rec, c.last_pair_count(), c.list_stats(), by_id(c.download(("id", "r", "v", "q", "L", "F", "M", "Vi"))))
    (rec, pairs, stats, sa), (rec0, pairs0, stats0, sb) = out["1"], out["0"]
    assert stats0["rebuilds"] == 0 and stats["rebuilds"] >= min_rebuilds, (stats0, stats)
    assert pairs == pairs0
    assert np.all(np.abs(rec - rec0) <= 1e-9 * np.maximum(np.abs(rec0).max(axis=0), 1e-300)), np.abs(rec - rec0).max(axis=0)
    for k in ("r", "v", "q", "L"):
        assert np.abs(sa[k] - sb[k]).max() <= 1e-9 * max(np.abs(sb[k]).max(), 1e-300), k
    for k in ("F", "M", "Vi"):
        assert vec_rel_err(sa[k], sb[k]) < 1e-6, k                       # 60 steps of trajectory divergence, not arithmetic


def test_multisite_group_list_loop_matches_single_rank():
    """CO2 on two in-process ranks: the halo refresh carries the orientations (7 doubles per halo molecule) next to the positions"""
    from tests.gpu_helpers import components_of
    d = load("ex_CO2")
    st = initial_state(d)
    box, rc, rclj, dt = d["box"], float(d["rc"][0]), float(d["rc"][1]), 4.0 * float(d["dt"][0])
    comps, mix = components_of(d), d["mixcoeff"]
    steps = 40
    with container_of(d) as c:
        c.add_particles(st["id"], st["cid"], st["r"], st["v"], st["q"], st["L"])
        c.exchange()
        c.traverse_cells()
        ref_rec = c.run(steps, dt)
        ref_stats = c.list_stats()
        ref = by_id(c.download(("id", "r", "v", "q", "L")))
    ranks = []
    try:
        for rank in range(2):
            grid, coords = ls1g.decomp_grid(2, rank)
            bmin, bmax = ls1g.decomp_box(grid, coords, box)
            c = ls1g.Container(bmax, rc, box_min=bmin, lj_cutoff=rclj, eps_rf=float(d["epsRF"][0]), global_length=box, rank_grid=grid,
                               rank_coords=coords, capacity=4 * len(st["id"]) + 1024, device=rank % ls1g.device_count())
            c.set_components(comps, mix)
            mine = np.all((st["r"] >= np.array(bmin)) & (st["r"] < np.array(bmax)), axis=1)
            c.add_particles(st["id"][mine], st["cid"][mine], st["r"][mine], st["v"][mine], st["q"][mine], st["L"][mine])
            ranks.append(c)
        ls1g.comm_init_local(ranks)
        ls1g.exchange_group(ranks)
        for c in ranks:
            c.traverse_cells()
        rec = ls1g.run_group(ranks, steps, dt)
        stats = ranks[0].list_stats()
        parts = [c.download(("id", "r", "v", "q", "L")) for c in ranks]
    finally:
        for c in ranks:
            c.close()
    out = by_id({k: np.concatenate([p[k] for p in parts]) for k in parts[0]})
    assert stats["rebuilds"] >= 2 and ref_stats["rebuilds"] >= 2
    assert np.array_equal(out["id"], ref["id"])
    assert np.all(np.abs(rec[:, :4] - ref_rec[:, :4]) <= 1e-9 * np.maximum(np.abs(ref_rec[:, :4]).max(axis=0), 1e-300))
    for k in ("r", "v", "q", "L"):
        assert np.abs(out[k] - ref[k]).max() <= 1e-9 * max(np.abs(ref[k]).max(), 1e-300), k


def test_exploded_system_is_an_error_not_a_fault():
    """a molecule that is still outside the halo region after the periodic wrap (it moved more than a box length in one step) ends the
    run with an error, as the reference's failing cell lookup does - no out-of-bounds access, the library stays usable"""
    r, v, box = fcc_lattice(6, seed=5)
    n = len(r)
    v = v.copy()
    v[7] = [4.0e4, 0.0, 0.0]
    with ls1g.Container(box, 2.5) as c:
        c.set_components([ls1g.lj_component()])
        c.add_particles(np.arange(n), np.zeros(n, dtype=np.uint32), r, v)
        c.exchange()
        c.traverse_cells()
        with pytest.raises(ls1g.Ls1gError, match="exploded"):
            c.run(3, 0.002)
    with ls1g.Container(box, 2.5) as c:                                   # a fresh container works as before
        c.set_components([ls1g.lj_component()])
        c.add_particles(np.arange(n), np.zeros(n, dtype=np.uint32), r, 0.0 * v)
        c.exchange()
        u, _ = c.traverse_cells()
        assert np.isfinite(u)


def test_group_thermostat_uses_the_global_kinetic_sums():
    """NVT on a local group: the betas come from the all-reduced sums (Domain::calculateGlobalValues collComm, Domain.cpp:170-176),
    so the decomposed run follows the single-rank one"""
    r, v, box = fcc_lattice(12, seed=3, temperature=1.0)
    n = len(r)
    ids = np.arange(n)
    with ls1g.Container(box, 2.5) as c:
        c.set_components([ls1g.lj_component()])
        c.add_particles(ids, np.zeros(n, dtype=np.uint32), r, v)
        c.exchange()
        c.traverse_cells()
        ref = c.run(30, 0.002, target_temperature=1.3)
    ranks = _group_of_ranks(4, box, 2.5, ids, r, v, capacity=4 * n + 4096)
    try:
        rec = ls1g.run_group(ranks, 30, 0.002, target_temperature=1.3)
    finally:
        for c in ranks:
            c.close()
    assert np.all(rec[:, 4] == n)
    assert np.all(np.abs(rec[:, :3] - ref[:, :3]) <= 1e-9 * np.abs(ref[:, :3]).max(axis=0))


@pytest.mark.parametrize("n_ranks", [2, 8])
@pytest.mark.parametrize("case", ["1clj", "ex_CO2"])
def test_rdf_decomposed_sums_to_single_rank(n_ranks, case):
    """block decomposition: every rank counts the pairs the reference counts on that rank (a pair across a rank boundary is seen from
    the side whose boundary cell index is below the halo cell index); summed over the ranks (RDF::collectRDF) the histograms are those
    of the single-rank traversal"""
    if case == "1clj":
        r, v, box = fcc_lattice(12, seed=16)
        rng = np.random.default_rng(18)
        r = np.mod(r + rng.normal(scale=0.05, size=r.shape), box)
        n = len(r)
        ids, cid, q, L = np.arange(n), np.zeros(n, dtype=np.uint32), None, None
        rc, rclj, comps, mix = 2.5, 2.5, [ls1g.lj_component()], ()
        shape = dict(bins=80, interval_length=0.03)
    else:
        d = load(case)
        st = initial_state(d)
        ids, cid, r, v, q, L = st["id"], st["cid"], st["r"], st["v"], st["q"], st["L"]
        box, rc, rclj = d["box"], float(d["rc"][0]), float(d["rc"][1])
        from tests.gpu_helpers import components_of
        comps, mix = components_of(d), d["mixcoeff"]
        n = len(ids)
        shape = dict(bins=40, interval_length=0.8, angular_bins=6)
    with ls1g.Container(box, rc, lj_cutoff=rclj) as c:
        c.set_components(comps, mix)
        c.add_particles(ids, cid, r, v, q, L)
        c.exchange()
        ref = c.rdf(**shape)
    ranks = []
    try:
        for rank in range(n_ranks):
            grid, coords = ls1g.decomp_grid(n_ranks, rank)
            bmin, bmax = ls1g.decomp_box(grid, coords, box)
            c = ls1g.Container(bmax, rc, box_min=bmin, lj_cutoff=rclj, global_length=box, rank_grid=grid,
                               rank_coords=coords, capacity=4 * n + 1024)
            c.set_components(comps, mix)
            mine = np.all((r >= np.array(bmin)) & (r < np.array(bmax)), axis=1)
            c.add_particles(ids[mine], cid[mine], r[mine], v[mine], None if q is None else q[mine], None if L is None else L[mine])
            ranks.append(c)
        ls1g.comm_init_local(ranks)
        ls1g.exchange_group(ranks)
        parts = [c.rdf(**shape) for c in ranks]
    finally:
        for c in ranks:
            c.close()
    assert ref["mol"].sum() > 10 * n
    for k in ref:
        assert np.array_equal(sum(p[k] for p in parts), ref[k]), k


@pytest.mark.parametrize("name", golden_cases("rdf_"))
def test_rdf_matches_reference(name):
    """ls1g_rdf (the reference's RDFCellProcessor + RDF counters as a device cell processor, adapter/RDFCellProcessor.cpp:18-133,
    io/RDF.h:95-160) against the reference's own counters after one traversal of the same cells - among them the inputs and
    expected counts of its RDFTest - and against the oracle restatement: molecule, site-site and angular histograms, every bin exact"""
    d = load(name)
    own = d["f_owned"].astype(bool)
    order = np.argsort(d["f_id"][own])
    n = int(own.sum())
    bins, interval, abins, _ = d["rdf_params"]
    with container_of(d) as c:
        c.add_particles(d["f_id"][own][order].astype(np.uint64), d["f_cid"][own][order].astype(np.uint32), d["f_r"][own][order],
                        np.zeros((n, 3)), d["f_q"][own][order], np.zeros((n, 3)))
        if d["meta"][0]:
            c.exchange()
        else:
            c.update()
        res = c.rdf(int(bins), float(interval), int(abins))
        again = c.rdf(int(bins), float(interval), int(abins))
    assert np.array_equal(res["mol"], d["rdf_mol"])
    assert np.array_equal(res["site"], d["rdf_site"])
    if abins > 1:
        assert np.array_equal(res["ardf"].ravel(), d["rdf_ardf"])
    for k in res:
        assert np.array_equal(res[k], again[k])
    s, g = system_and_grid(d)
    ref = O.rdf(s, g, d["f_cid"], d["f_r"], d["f_q"], int(bins), float(interval), int(abins))
    assert np.array_equal(res["mol"], ref["mol"]) and np.array_equal(res["site"], ref["site"])


def test_rdf_of_the_bench_lattice_after_steps():
    """fcc 1CLJ (bench.py's workload, 16^3 cells): ls1g_rdf after 10 device-resident steps (leaves the neighbour-list layout, halo copies
    rebuilt) equals the oracle on the downloaded state; the forces survive the re-sort (the next step continues the same trajectory)"""
    r, v, box = fcc_lattice(16, seed=21)
    ids = np.arange(len(r))

    def fresh():
        c = ls1g.Container(box, 2.5)
        c.set_components([ls1g.lj_component()])
        c.add_particles(ids, np.zeros(len(ids), dtype=np.uint32), r, v)
        c.exchange()
        c.traverse_cells()
        return c

    with fresh() as c, fresh() as plain:
        c.run(10, 0.002)
        plain.run(10, 0.002)
        c.exchange()
        res = c.rdf(100, 0.025)
        allm = c.download(("id", "cid", "r"), include_halo=True)
        c.delete_outer_particles()
        a, b = c.run(5, 0.002), plain.run(5, 0.002)
    assert np.allclose(a, b, rtol=1e-10, atol=0)
    s = O.System.single_lj(rc=2.5)
    g = O.grid([0, 0, 0], box, 2.5, 0)
    q = np.tile(np.array([1.0, 0, 0, 0]), (len(allm["id"]), 1))
    ref = O.rdf(s, g, allm["cid"], allm["r"], q, 100, 0.025)
    assert res["mol"].sum() > 20 * len(ids) and np.array_equal(res["mol"], ref["mol"])


def test_reserve_allocates_ahead_of_the_upload():
    """ls1g_reserve: the device arrays exist before the first upload (set-up cost outside a timed upload); results are unchanged and a
    loaded context refuses to be re-allocated"""
    r, v, box = fcc_lattice(6, seed=2)
    n = len(r)
    out = []
    for reserve in (False, True):
        with ls1g.Container(box, 2.5) as c:
            c.set_components([ls1g.lj_component()])
            if reserve:
                c.reserve(n)
                c.reserve(n)
            c.add_particles(np.arange(n), np.zeros(n, dtype=np.uint32), r, v)
            c.exchange()
            u, w = c.traverse_cells()
            out.append((u, w, by_id(c.download(("id", "F")))["F"]))
            if reserve:
                with pytest.raises(ls1g.Ls1gError, match="already holds molecules"):
                    c.reserve(100 * n)
    assert out[0][0] == out[1][0] and out[0][1] == out[1][1] and np.array_equal(out[0][2], out[1][2])


@pytest.mark.parametrize("name", ["ex_CO2", "per_LennardJones"])
def test_slot_ranges_hold_the_owned_molecules_in_download_order(name):
    """ls1g_download_range: raw slots of the cell-sorted storage, chunk by chunk; the slots without HALO_FLAG are exactly what
    ls1g_download returns, in the same order - before and after the halo copies are dropped"""
    d = load(name)
    with container_of(d) as c:
        _load_into(c, d, True)
        c.traverse_cells()
        for dropped in (False, True):
            if dropped:
                c.delete_outer_particles()
            ref = c.download(("id", "cid", "r", "v", "q", "L", "F", "M", "Vi", "cell"))
            full = c.download_slots(("cell",))
            ns = len(full["cell"])
            assert ns > len(ref["id"])
            parts = [c.download_slots(("id", "cid", "r", "v", "q", "L", "F", "M", "Vi", "cell"), a, min(ns, a + 700) - a) for a in range(0, ns, 700)]
            got = {k: np.concatenate([p[k] for p in parts]) for k in parts[0]}
            assert np.array_equal(got["cell"], full["cell"])
            own = (got["cell"] & ls1g.binding.HALO_FLAG) == 0
            assert own.sum() == len(ref["id"])
            for k in ref:
                assert np.array_equal(got[k][own], ref[k]), k
        with pytest.raises(ls1g.Ls1gError, match="range outside"):
            c.download_slots(("id",), ns - 1, 2)
