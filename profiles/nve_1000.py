"""BASELINE.json configs[1]: synthetic 1CLJ fcc lattice, 10^7 molecules, 1 B200, NVE 1000 steps - energy drift check.
Runs the device-resident loop (ls1g_run) for 1000 steps with the unshifted LJ potential of cfg B and with the shifted one
(Component::addLJcenter shift6, Component.cpp:105-117: the unshifted potential jumps by U(rc) whenever a pair crosses the cutoff,
so only the shifted energy is a conserved quantity) and prints one JSON line.  tests/test_gpu_parity.py::test_nve_drift_matches_oracle
holds the same loop to the CPU oracle's energy series (1e-8) at a size the oracle finishes in seconds.
Usage (GPU box): python profiles/nve_1000.py [cells=136] [steps=1000]"""
import importlib, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
ls1g = importlib.import_module("ls1-mardyn_b200")
gen = importlib.import_module("ls1-mardyn_b200.generators")
cells = int(sys.argv[1]) if len(sys.argv) > 1 else 136
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
RHO, TEMP, RC, DT = 0.8442, 1.44, 2.5, 0.002
ids, r, v = gen.fcc_block(cells, rho=RHO, temperature=TEMP, seed=1)
v -= v.mean(axis=0)
_, L = gen.fcc_box(cells, RHO)
n = len(ids)
out = {"workload": "1CLJ fcc rho*=0.8442 T*=1.44 rc=2.5 dt*=0.002 NVE, %d molecules, %d steps" % (n, steps)}
s6 = (1.0 / RC) ** 6
for name, shift6 in (("unshifted", 0.0), ("shifted", 24.0 * (s6 - s6 * s6))):
    with ls1g.Container(list(L), RC, capacity=int(1.35 * n) + 200000) as c:
        c.set_components([ls1g.lj_component(shift6=shift6)])
        c.add_particles(ids, np.zeros(n, dtype=np.uint32), r, v)
        c.exchange()
        u0, _ = c.traverse_cells()
        t0 = time.perf_counter()
        rec = c.run(steps, DT)
        wall = time.perf_counter() - t0
    e = rec[:, 0] + 0.5 * rec[:, 2]
    ekin = 0.5 * rec[:, 2]
    out[name] = {"e_total_first": e[0], "e_total_last": e[-1], "rel_drift_last_vs_first": abs(e[-1] - e[0]) / abs(e[0]),
                 "max_rel_deviation": float(np.abs(e - e[0]).max() / abs(e[0])), "drift_over_mean_kinetic_energy": float(abs(e[-1] - e[0]) / ekin.mean()),
                 "T_first": rec[0, 2] / (3 * n), "T_last": rec[-1, 2] / (3 * n), "n_constant": bool(np.all(rec[:, 4] == n)),
                 "mmups": n * steps / wall / 1e6, "wall_s": wall}
print(json.dumps(out))
