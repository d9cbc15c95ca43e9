"""Device time of ls1g_rdf (csrc/rdf.cu) on bench.py's workload and on replicated CO2 (site-site + angular tables).
Run on the GPU box:  python profiles/rdf_timing.py [cells] > gpurun_out/rdf_timing.json
CUDA-event time of the kernels (LS1G_T_AUX timer: rdf_sites_kernel + rdf_kernel), not the host call."""
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import bench_examples as BE  # noqa: E402

ls1g = importlib.import_module("ls1-mardyn_b200")


def timed(c, n, **kw):
    c.rdf(**kw)                                   # warm-up (function attributes, allocation)
    c.timers(reset=True)
    t0 = time.perf_counter()
    res = [c.rdf(**kw) for _ in range(3)]
    wall = (time.perf_counter() - t0) / 3
    t = c.timers()["aux"]
    pairs = int(res[0]["mol"].sum())
    fc = c.flop_count()
    ms = t["ms"] / 3
    return dict(n_molecules=int(n), kernel_ms=round(ms, 3), host_call_ms=round(wall * 1e3, 3), launches_per_call=t["launches"] // 3,
                pairs_binned=pairs, site_pairs_binned=int(res[0]["site"].sum()), ardf_entries=int(res[0]["ardf"].sum()),
                molecule_distances=fc["molecule_distances"],
                candidates_per_s=round(fc["molecule_distances"] / (ms * 1e-3), 0),
                algorithmic_GBps=round(fc["molecule_distances"] * 32 / (ms * 1e-3) / 1e9, 1), **kw)


def main():
    cells = int(sys.argv[1]) if len(sys.argv) > 1 else 136
    out = {}
    sh = bench.make_shard([cells] * 3, 1, 0)
    with ls1g.Container(sh["L"], bench.RC) as c:
        c.set_components([ls1g.lj_component()])
        c.add_particles(sh["id"], np.zeros(len(sh["id"]), dtype=np.uint32), sh["r"], sh["v"])
        c.exchange()
        c.traverse_cells()
        c.run(20, bench.DT)                      # a liquid-like state instead of the perfect lattice
        c.exchange()
        out["1clj_fcc"] = timed(c, len(sh["id"]), bins=100, interval_length=0.025)
    d = BE.fixture("ex_CO2")
    m, box = BE.replicate(d, 8)                   # 2048 x 8^3 = 1 048 576 molecules, 3 LJ + 3 charges each
    comps = ls1g.components_from_tables(d["comp_counts"], d["comp_scalars"], d["lj_sites"], d["charge_sites"], d["dipole_sites"], d["quadrupole_sites"])
    with ls1g.Container(box, float(d["rc"][0]), lj_cutoff=float(d["rc"][1]), eps_rf=float(d["epsRF"][0])) as c:
        c.set_components(comps, d["mixcoeff"])
        c.add_particles(m["id"], m["cid"], m["r"], m["v"], q=m["q"], L=m["D"])
        c.exchange()
        out["co2_x8"] = timed(c, len(m["id"]), bins=50, interval_length=0.7, angular_bins=8)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
