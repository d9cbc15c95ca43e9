"""Throughput of the multi-site path on the reference's shipped examples (SURVEY 8 cfg A/C/D), replicated k^3 like the
reference's ReplicaGenerator: python profiles/probe_multisite.py ex_CO2 8 20   (fixture, replicas per dim, steps)."""
import importlib, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
ls1g = importlib.import_module("ls1-mardyn_b200")


def replicate(d, k):
    box = d["box"]
    sh = np.array([(i, j, l) for i in range(k) for j in range(k) for l in range(k)], dtype=np.float64) * box
    n = len(d["in_id"])
    r = (d["in_r"][None, :, :] + sh[:, None, :]).reshape(-1, 3)
    rep = lambda a: np.tile(a, (k ** 3,) + (1,) * (a.ndim - 1))
    return dict(id=np.arange(n * k ** 3, dtype=np.uint64), cid=rep(d["in_cid"]).astype(np.uint32), r=r, v=rep(d["in_v"]),
                q=rep(d["in_q"]), L=rep(d["in_L"])), box * k


def main():
    name, k, steps = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
    d = np.load(os.path.join(ROOT, "tests", "golden", name + ".npz"))
    m, box = replicate(d, k)
    comps = ls1g.components_from_tables(d["comp_counts"], d["comp_scalars"], d["lj_sites"], d["charge_sites"], d["dipole_sites"], d["quadrupole_sites"])
    with ls1g.Container(box, float(d["rc"][0]), lj_cutoff=float(d["rc"][1]), eps_rf=float(d["epsRF"][0])) as c:
        c.set_components(comps, d["mixcoeff"])
        c.add_particles(m["id"], m["cid"], m["r"], m["v"], q=m["q"], L=m["L"])
        c.exchange()
        u, w = c.traverse_cells()
        dt = float(d["dt"][0])
        c.run(3, dt)
        c.timers(reset=True)
        t0 = time.perf_counter()
        c.run(steps, dt)
        t = time.perf_counter() - t0
        tm = c.timers(reset=True)
        n = len(m["id"])
        print(name, "N", n, "steps", steps, "ms/step %.3f" % (t / steps * 1e3), "MMUPS %.1f" % (n * steps / t / 1e6),
              {kk: round(v["ms"] / steps, 3) for kk, v in tm.items()}, "upot/N %.6g (k=1 ref %.6g)" % (u / n, d["upot_virial"][0] / len(d["in_id"])),
              "ref flops/molecule %.0f" % (d["flops"][0] / len(d["in_id"])))


if __name__ == "__main__":
    main()
