"""Times the pieces of the host<->device path of one e2e step (upload, forces, run(1), download) and the raw pinned copy rate.
Usage (GPU box): python profiles/probe_transfer.py [cells]"""
import importlib, os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
ls1g = importlib.import_module("ls1-mardyn_b200")
gen = importlib.import_module("ls1-mardyn_b200.generators")
cells = int(sys.argv[1]) if len(sys.argv) > 1 else 136
ids, r, v = gen.fcc_block(cells, rho=0.8442, temperature=1.44, seed=1)
_, L = gen.fcc_box(cells, 0.8442)
n = len(ids); cap = int(1.35 * n) + 200000
def pinned(shape, dt): return torch.empty(shape, dtype=dt, pin_memory=True).numpy()
host = dict(id=pinned((cap,), torch.int64).view(np.uint64), cid=pinned((cap,), torch.int32).view(np.uint32),
            r=pinned((cap, 3), torch.float64), v=pinned((cap, 3), torch.float64), F=pinned((cap, 3), torch.float64))
host["id"][:n] = ids; host["cid"][:] = 0; host["r"][:n] = r; host["v"][:n] = v
# raw pinned copy rate
t = torch.from_numpy(host["r"]); d = torch.empty_like(t, device="cuda")
for _ in range(2):
    torch.cuda.synchronize(); t0 = time.perf_counter(); d.copy_(t, non_blocking=True); torch.cuda.synchronize(); dt = time.perf_counter() - t0
print("raw H2D pinned %.1f GB/s" % (t.numel() * 8 / dt / 1e9))
for _ in range(2):
    torch.cuda.synchronize(); t0 = time.perf_counter(); t.copy_(d, non_blocking=True); torch.cuda.synchronize(); dt = time.perf_counter() - t0
print("raw D2H pinned %.1f GB/s" % (t.numel() * 8 / dt / 1e9))
del d
c = ls1g.Container(list(L), 2.5, capacity=cap)
c.set_components([ls1g.lj_component()])
c.add_particles(host["id"][:n], host["cid"][:n], host["r"][:n], host["v"][:n])
c.exchange(); c.traverse_cells(); c.run(3, 0.002)
c.delete_outer_particles()
m = len(c.download(("id", "r", "v", "F"), out=host)["id"])
for it in range(3):
    T = []
    t0 = time.perf_counter(); c.add_particles(host["id"][:m], host["cid"][:m], host["r"][:m], host["v"][:m]); T.append(time.perf_counter() - t0)
    t0 = time.perf_counter(); c.set_forces(host["F"][:m]); T.append(time.perf_counter() - t0)
    t0 = time.perf_counter(); c.run(1, 0.002, records=True); T.append(time.perf_counter() - t0)
    t0 = time.perf_counter(); m = len(c.download(("id", "r", "v", "F"), out=host)["id"]); T.append(time.perf_counter() - t0)
    print("iter %d: upload %.1f ms, forces %.1f ms, run(1) %.1f ms, download %.1f ms" % (it, *[1e3 * x for x in T]))
t0 = time.perf_counter(); c.run(1, 0.002, records=True); print("run(1) again %.1f ms" % (1e3 * (time.perf_counter() - t0)))
t0 = time.perf_counter(); c.run(1, 0.002, records=True); print("run(1) again %.1f ms" % (1e3 * (time.perf_counter() - t0)))
c.close()
